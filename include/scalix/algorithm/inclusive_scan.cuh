// scalix-b200: algorithm::inclusive_scan (surface of /root/reference
// scalix/algorithm/inclusive_scan.cuh:44-207, four overloads).
//
// Scans the FLATTENED storage order whatever the rank, accumulating in the input's value type;
// `identity` is accepted and ignored, exactly like the reference (ibid. :48,87-93).
//
//   reference : per device thrust::inclusive_scan on a fresh stream from a fresh thread, then a
//               serial HOST carry chain through Unified Memory, then a second full fix-up pass
//               over every partition but the first (ibid. :75-157).
//   here      : 1 partition  -> ONE streaming scan kernel (scan_stream.cuh): every element is
//                               read once and written once.
//               N partitions -> pass A: per-partition totals (reduce_kernel, reads each element
//                               once) written to peer-visible scratch; pass B: the same scan
//                               kernel on every partition, seeded on the device with the chained
//                               totals of the partitions below it.  No host round trip, no
//                               fix-up pass; A and B are issued in one fork-join and ordered
//                               across devices by stream events.
//                               This path needs an associative AND commutative operator with a
//                               known neutral element (the thrust functors of functional.cuh);
//                               any other functor uses the reference's scan + chain + fix-up
//                               scheme on top of the same scan kernel.
// Operand order is the reference's: op(earlier, later) inside a partition; the carry is applied
// as f(x, carry), the partition's last element is f(carry_chain, last) (ibid. :122,150).
#pragma once
#include "../assign_array.cuh"
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"
#include "functional.cuh"

#include <vector>

namespace sclx::algorithm {

namespace detail {

// x = f(x, carry) for every element of a partition except its last one, which becomes
// `last_value` (the chained total): the fix-up of scalix/algorithm/inclusive_scan.cuh:125-157
template<class T, class F>
__global__ void scan_fixup_kernel(T* data, size_t n, T carry, T last_value, F f) {
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        data[i] = (i + 1 == n) ? last_value : f(data[i], carry);
}

template<class T>
constexpr bool scan_kernel_supports_v
    = std::is_trivially_copyable_v<T> && (sizeof(T) == 1 || sizeof(T) == 2 || sizeof(T) == 4 || sizeof(T) == 8);

}  // namespace detail

template<class T, class ResultT, uint Rank, class F>
__host__ void inclusive_scan(
    const array<const T, Rank>& arr,
    array<ResultT, Rank>& result,
    const T& identity,
    F&& f
) {
    if (arr.shape() != result.shape()) {
        throw_exception<std::invalid_argument>(
            "Input and result arrays must have the same shape", "sclx::algorithm::"
        );
    }
    (void)identity;
    using functor_t = std::decay_t<F>;
    static_assert(
        detail::scan_kernel_supports_v<T>,
        "inclusive_scan: element types must be trivially copyable and 1, 2, 4 or 8 bytes wide"
    );

    if constexpr (!std::is_same_v<T, ResultT>) {
        // accumulate in T like the reference, then convert
        array<T, Rank> tmp(arr.shape());
        inclusive_scan(arr, tmp, identity, f);
        assign_array(tmp, result).get();
        return;
    } else {
        const int caller_device = cuda::traits::current_device();
        const auto device_info  = get_device_split_info(arr);
        const int nparts        = int(device_info.size());
        const T* in             = arr.data().get();
        T* out                  = result.data().get();
        functor_t op(f);
        b200::call_workspaces ws(device_info);  // leased for this call only

        if (nparts == 1) {
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                b200::launch_inclusive_scan<T, functor_t>(
                    in + p.offset, out + p.offset, p.elements, nullptr, 0, nullptr, op, p, ws
                );
            });
        } else if constexpr (detail::is_known_commutative_v<functor_t> && std::is_arithmetic_v<T>) {
            // pass A: totals of every partition but the last, into peer-visible scratch
            T* totals       = ws.scratch<T>(size_t(nparts));
            const T neutral = detail::neutral_element<functor_t, T>();
            // pass A (totals of every partition but the last) and pass B (the scan, seeded with
            // the chained totals of the partitions below) in ONE fork-join: device d's stream
            // waits for pass A of the devices below it between its two kernels
            b200::for_each_partition_chained(
                arr, device_info, /*lower_only=*/true,
                [&](const b200::partition_t& p) {
                    if (p.ordinal + 1 < nparts)
                        b200::launch_reduce<T, functor_t>(
                            in + p.offset, p.elements, neutral, totals + p.ordinal, op, p, ws
                        );
                },
                [&](const b200::partition_t& p) {
                    b200::launch_inclusive_scan<T, functor_t>(
                        in + p.offset, out + p.offset, p.elements, totals, p.ordinal, nullptr, op, p, ws
                    );
                }
            );
        } else {
            // generic functor: the reference's three stages on top of the scan kernel
            T* lasts = ws.scratch<T>(size_t(nparts));
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                b200::launch_inclusive_scan<T, functor_t>(
                    in + p.offset, out + p.offset, p.elements, nullptr, 0, lasts + p.ordinal, op, p, ws
                );
            });
            const auto devices = b200::devices_of(device_info);
            b200::sync_devices(devices);
            std::vector<T> chain(static_cast<size_t>(nparts), T{});
            cuda::cuda_exception::raise_if_not_success(
                cudaMemcpy(chain.data(), lasts, sizeof(T) * size_t(nparts), cudaMemcpyDeviceToHost),
                std::experimental::source_location::current(),
                "sclx::algorithm::"
            );
            std::vector<T> carry(static_cast<size_t>(nparts), T{});  // carry[d] = chained total of partitions < d
            for (int d = 1; d < nparts; ++d) {
                carry[d] = chain[d - 1];
                chain[d] = op(chain[d - 1], chain[d]);
            }
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                if (p.ordinal == 0 || p.elements == 0) return;
                const size_t work = (p.elements + 255) / 256;
                detail::scan_fixup_kernel<T, functor_t>
                    <<<b200::persistent_grid(work, 8, b200::sm_count_of(p.device)), 256, 0, p.stream>>>(
                        out + p.offset, p.elements, carry[p.ordinal], chain[p.ordinal], op
                    );
            });
        }
        b200::sync_devices(b200::devices_of(device_info));
        cuda::set_device(caller_device);
    }
}

template<class T, class ResultT, uint Rank, class F>
__host__ void inclusive_scan(
    const array<T, Rank>& arr,
    array<ResultT, Rank>& result,
    const T& identity,
    F&& f
) {
    inclusive_scan(static_cast<const array<const T, Rank>&>(arr), result, identity, f);
}

template<class T, uint Rank, class F>
__host__ array<T, Rank>
inclusive_scan(const array<const T, Rank>& arr, const T& identity, F&& f) {
    array<T, Rank> result(arr.shape());
    inclusive_scan(arr, result, identity, f);
    return result;
}

template<class T, uint Rank, class F>
__host__ array<T, Rank> inclusive_scan(const array<T, Rank>& arr, const T& identity, F&& f) {
    array<T, Rank> result(arr.shape());
    inclusive_scan(static_cast<const array<const T, Rank>&>(arr), result, identity, f);
    return result;
}

}  // namespace sclx::algorithm
