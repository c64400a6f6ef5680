// scalix-b200: implementation detail of array_traits.cuh
// (surface of /root/reference scalix/detail/array_traits.cuh:35-43).
#pragma once
#include "../array.cuh"
#include <type_traits>

namespace sclx::detail {

template<class>
struct is_scalix_array : std::false_type {};

template<class T, uint Rank>
struct is_scalix_array<array<T, Rank>> : std::true_type {};

}  // namespace sclx::detail
