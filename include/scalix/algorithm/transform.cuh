// scalix-b200: algorithm::transform (surface of /root/reference
// scalix/algorithm/transform.cuh:41-64): result[i] = op(arr[i], scalar), asynchronous.
#pragma once
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"
#include "../throw_exception.hpp"
#include "functional.cuh"

namespace sclx::algorithm {

REGISTER_SCALIX_KERNEL_TAG(transform_kernel);  // kept for source compatibility

template<class R, class T, uint Rank, class U, class BinaryOp>
std::future<void> transform(
    sclx::array<R, Rank>& result,
    const sclx::array<T, Rank>& arr,
    const U& scalar,
    BinaryOp&& op
) {
    if (result.shape() != arr.shape()) {
        throw_exception<std::invalid_argument>("input shapes must match", "sclx::algorithm::");
    }
    return sclx::execute_kernel([&](const kernel_handler&) {
        const T* src = arr.data().get();
        R* dst       = result.data().get();
        std::decay_t<BinaryOp> fn(op);
        const U value = scalar;
        b200::for_each_partition(result, [&](const b200::partition_t& p) {
            b200::launch_generate(
                dst + p.offset, p.elements, p.offset,
                [=] __device__(size_t i) { return static_cast<R>(fn(src[i], value)); }, p
            );
        });
    });
}

}  // namespace sclx::algorithm
