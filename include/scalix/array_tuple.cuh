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

// The tuple is a thrust::tuple so that it is trivially copyable into a kernel's parameter
// space together with the device lambda; the launcher takes its partitioning from element 0
// and requires equal last extents of all elements (kernel_handler::launch_tuple_impl).
// Element types may differ (array<float,1> next to array<int,3>), which is what array_list
// cannot express.
template<class... Types>
class array_tuple : public thrust::tuple<Types...> {
  public:
    using thrust::tuple<Types...>::tuple;

    static_assert(
        is_scalix_arrays_v<Types...>,
        "array_tuple must be constructed from scalix arrays"
    );

    // true when every array's slowest (partitioned) extent is the same — the precondition of
    // a multi-result launch
    __host__ __device__ bool last_dims_all_equal() {
        return apply(
            [](const auto&... arrays) { return detail::is_last_dim_of_arrays_equal(arrays...); },
            *this
        );
    }

    // Unified-Memory hints of the reference, forwarded to every element.  Arrays here live in
    // explicit per-device pages, so these are accepted and carry no effect (array.cuh).
    __host__ void unset_read_mostly() {
        apply([](auto&&... arrays) { (arrays.unset_read_mostly(), ...); }, *this);
    }

    __host__ void set_read_mostly() {
        apply([](auto&&... arrays) { (arrays.set_read_mostly(), ...); }, *this);
    }
};

// f(arrays...) with the tuple's elements expanded into an argument pack (host and device)
template<class F, class... Types>
__host__ __device__ auto apply(F&& f, array_tuple<Types...> t) {
    return detail::apply<F, sizeof...(Types), array_tuple<Types...>>(std::forward<F>(f), t);
}

// make_array_tuple(a, b, c): deduces the element types, e.g. for
//     handler.launch(range, sclx::make_array_tuple(out_f, out_i), lambda);
template<class... Types>
array_tuple<Types...> make_array_tuple(Types&&... args) {
    return array_tuple<Types...>(std::forward<Types>(args)...);
}

}  // namespace sclx
