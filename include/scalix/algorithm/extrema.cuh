// scalix-b200: algorithm::min_element / max_element (surface of /root/reference
// scalix/algorithm/extrema.cuh:44-159): the extreme VALUE of the whole array as a host scalar.
// Same kernel family as algorithm::reduce with thrust::minimum / thrust::maximum.
#pragma once
#include "reduce.cuh"

#include <limits>

namespace sclx::algorithm {

template<class T, uint Rank>
__host__ T min_element(const array<T, Rank>& arr) {
    using value_t = std::remove_const_t<T>;
    return reduce(arr, std::numeric_limits<value_t>::max(), minimum<>());
}

template<class T, uint Rank>
__host__ T max_element(const array<T, Rank>& arr) {
    using value_t = std::remove_const_t<T>;
    return reduce(arr, std::numeric_limits<value_t>::lowest(), maximum<>());
}

}  // namespace sclx::algorithm
