// scalix-b200 umbrella header: like /root/reference scalix/scalix.cuh:34-35 it only pulls in
// execute_kernel.cuh; algorithms are included individually.
#pragma once
#include "execute_kernel.cuh"
