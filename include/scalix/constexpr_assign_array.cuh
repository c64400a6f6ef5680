// scalix-b200: compile-time friendly element copy (API of
// /root/reference scalix/constexpr_assign_array.cuh:37-44).
#pragma once

namespace sclx {

template<unsigned N, class T>
__host__ __device__ constexpr void constexpr_assign_array(T* dest, const T* src) {
    for (unsigned i = 0; i < N; ++i) dest[i] = src[i];
}

}  // namespace sclx
