// scalix-b200: type traits for sclx::array
// (surface of /root/reference scalix/array_traits.cuh:39-58).
#pragma once
#include "detail/array_traits.cuh"

namespace sclx {

template<class T>
struct is_scalix_array : detail::is_scalix_array<std::decay_t<T>> {};

template<class T>
constexpr bool is_scalix_array_v = is_scalix_array<T>::value;

template<class... Types>
struct is_scalix_arrays : std::conjunction<is_scalix_array<Types>...> {};

template<class T, class... Types>
constexpr bool is_scalix_arrays_v = is_scalix_arrays<T, Types...>::value;

}  // namespace sclx
