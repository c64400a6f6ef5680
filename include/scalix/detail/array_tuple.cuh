// scalix-b200: helpers of array_tuple.cuh
// (surface of /root/reference scalix/detail/array_tuple.cuh:38-77).
#pragma once
#include "../array_traits.cuh"
#include <thrust/tuple.h>
#include <utility>

namespace sclx::detail {

template<class... Types>
__host__ __device__ bool is_last_dim_of_arrays_equal(const Types&... arrays) {
    static_assert(
        is_scalix_arrays_v<Types...>,
        "is_last_dim_of_arrays_equal must be called with scalix arrays"
    );
    const size_t dims[] = {arrays.shape()[arrays.rank() - 1]...};
    for (size_t i = 1; i < sizeof...(Types); ++i)
        if (dims[i] != dims[0]) return false;
    return true;
}

template<class F, class Tuple, size_t... I>
__host__ __device__ auto apply_impl(F&& f, Tuple& t, std::index_sequence<I...>) {
    return f(thrust::get<I>(t)...);
}

// f(get<0>(t), ..., get<N-1>(t)); extra leading template arguments mirror the reference
template<class F, uint N, class Tuple>
__host__ __device__ auto apply(F&& f, Tuple t) {
    return apply_impl(std::forward<F>(f), t, std::make_index_sequence<N>{});
}

}  // namespace sclx::detail
