// scalix-b200: fill / zeros / ones (surface of /root/reference scalix/fill.cuh:40-79).
// The reference runs `arr[index] = value` through the per-element lambda launcher (one scalar
// store per thread, 64-bit div/mod per element for Rank > 1).  The access pattern is known, so
// here every partition is filled by fill_kernel with 128-bit stores; blocking, like the
// reference's `.get()`.
#pragma once
#include "detail/b200/host_launch.cuh"
#include "execute_kernel.cuh"

namespace sclx {

template<class T, uint Rank>
void fill(const array<T, Rank>& arr, const T& value) {
    if constexpr (std::is_trivially_copyable_v<T>) {
        const auto device_info = get_device_split_info(arr);
        T* base                = arr.data().get();
        b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
            b200::launch_fill(base + p.offset, p.elements, value, p);
        });
        b200::sync_devices(b200::devices_of(device_info));
    } else {
        execute_kernel([&](kernel_handler& handler) {
            handler.launch(
                md_range_t<Rank>(arr.shape()),
                arr,
                [=] __device__(const md_index_t<Rank>& index, const auto&) { arr[index] = value; }
            );
        }).get();
    }
}

template<class T, uint Rank>
__host__ array<T, Rank> zeros(const size_t (&shape)[Rank]) {
    // fresh allocations are zero-filled by the runtime
    return array<T, Rank>{shape_t<Rank>(shape)};
}

template<class T, uint Rank>
__host__ array<T, Rank> zeros(const shape_t<Rank>& shape) {
    return array<T, Rank>{shape};
}

template<class T, uint Rank>
__host__ array<T, Rank> ones(const size_t (&shape)[Rank]) {
    array<T, Rank> arr{shape_t<Rank>(shape)};
    fill(arr, T(1));
    return arr;
}

template<class T, uint Rank>
__host__ array<T, Rank> ones(const shape_t<Rank>& shape) {
    array<T, Rank> arr{shape};
    fill(arr, T(1));
    return arr;
}

}  // namespace sclx
