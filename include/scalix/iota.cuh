// scalix-b200: iota (surface of /root/reference scalix/iota.cuh:40-51): arr[i] = T(start) + i
// for a rank-1 array, blocking.  One iota_kernel launch per partition (128-bit stores), each
// writing its own slice of the one global sequence.
#pragma once
#include "detail/b200/host_launch.cuh"
#include "execute_kernel.cuh"

namespace sclx {

template<class T, class U>
__host__ void iota(sclx::array<T, 1>& arr, const U& start = 0) {
    if constexpr (std::is_arithmetic_v<T>) {
        const auto device_info = get_device_split_info(arr);
        T* base                = arr.data().get();
        const T first          = static_cast<T>(start);
        b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
            b200::launch_iota(base + p.offset, p.elements, first, p.offset, p);
        });
        b200::sync_devices(b200::devices_of(device_info));
    } else {
        sclx::execute_kernel([&](const sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{arr.shape()},
                arr,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    arr[idx] = static_cast<T>(start) + idx[0];
                }
            );
        }).get();
    }
}

}  // namespace sclx
