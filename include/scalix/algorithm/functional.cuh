// scalix-b200: functor aliases (surface of /root/reference scalix/algorithm/functional.cuh:38-60)
// plus the traits the algorithms use to pick their fast paths.  The aliases stay Thrust's
// functors so user code and fast-path detection see the same types as with the reference.
#pragma once
#include <thrust/functional.h>

#include <limits>
#include <type_traits>

namespace sclx::algorithm {

template<class T = void>
using plus = thrust::plus<T>;
template<class T = void>
using minus = thrust::minus<T>;
template<class T = void>
using multiplies = thrust::multiplies<T>;
template<class T = void>
using divides = thrust::divides<T>;
template<class T = void>
using maximum = thrust::maximum<T>;
template<class T = void>
using minimum = thrust::minimum<T>;
template<class T = void>
using max = thrust::maximum<T>;
template<class T = void>
using min = thrust::minimum<T>;

namespace detail {

enum class known_op { none, plus, multiplies, maximum, minimum };

template<class F>
struct classify : std::integral_constant<known_op, known_op::none> {};
template<class T>
struct classify<thrust::plus<T>> : std::integral_constant<known_op, known_op::plus> {};
template<class T>
struct classify<thrust::multiplies<T>> : std::integral_constant<known_op, known_op::multiplies> {};
template<class T>
struct classify<thrust::maximum<T>> : std::integral_constant<known_op, known_op::maximum> {};
template<class T>
struct classify<thrust::minimum<T>> : std::integral_constant<known_op, known_op::minimum> {};

// associative AND commutative with a known neutral element
template<class F>
constexpr bool is_known_commutative_v = classify<std::decay_t<F>>::value != known_op::none;

template<class F, class T>
constexpr T neutral_element() {
    constexpr known_op op = classify<std::decay_t<F>>::value;
    if constexpr (op == known_op::plus) return T(0);
    else if constexpr (op == known_op::multiplies) return T(1);
    else if constexpr (op == known_op::maximum) return std::numeric_limits<T>::lowest();
    else return std::numeric_limits<T>::max();
}

}  // namespace detail

}  // namespace sclx::algorithm
