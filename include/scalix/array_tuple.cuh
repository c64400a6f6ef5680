// scalix-b200: heterogeneous tuple of arrays for multi-result launches
// (surface of /root/reference scalix/array_tuple.cuh:45-81).
#pragma once
#include "array_traits.cuh"
#include "detail/array_tuple.cuh"

namespace sclx {

template<class... Types>
class array_tuple;

template<class F, class... Types>
__host__ __device__ auto apply(F&& f, array_tuple<Types...> t);

template<class... Types>
class array_tuple : public thrust::tuple<Types...> {
  public:
    using thrust::tuple<Types...>::tuple;

    static_assert(
        is_scalix_arrays_v<Types...>,
        "array_tuple must be constructed from scalix arrays"
    );

    __host__ __device__ bool last_dims_all_equal() {
        return apply(
            [](const auto&... arrays) { return detail::is_last_dim_of_arrays_equal(arrays...); },
            *this
        );
    }

    __host__ void unset_read_mostly() {
        apply([](auto&&... arrays) { (arrays.unset_read_mostly(), ...); }, *this);
    }

    __host__ void set_read_mostly() {
        apply([](auto&&... arrays) { (arrays.set_read_mostly(), ...); }, *this);
    }
};

template<class F, class... Types>
__host__ __device__ auto apply(F&& f, array_tuple<Types...> t) {
    return detail::apply<F, sizeof...(Types), array_tuple<Types...>>(std::forward<F>(f), t);
}

template<class... Types>
array_tuple<Types...> make_array_tuple(Types&&... args) {
    return array_tuple<Types...>(std::forward<Types>(args)...);
}

}  // namespace sclx
