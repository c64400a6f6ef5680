// scalix-b200: algorithm::reduce (surface of /root/reference scalix/algorithm/reduce.cuh:44-101):
// whole-array fold to a host scalar.  One reduce_kernel per partition (128-bit loads, one pass),
// partials in peer-visible scratch, folded on the host in FIXED partition order (the reference
// folds in mutex-acquisition order).  `f` must be associative and commutative.
#pragma once
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"
#include "functional.cuh"

#include <vector>

namespace sclx::algorithm {

template<class T, uint Rank, class F>
__host__ T reduce(const array<T, Rank>& arr, const T& identity, F&& f) {
    using value_t   = std::remove_const_t<T>;
    using functor_t = std::decay_t<F>;
    static_assert(std::is_trivially_copyable_v<value_t>, "reduce: trivially copyable elements only");
    const int caller_device = cuda::traits::current_device();
    const auto device_info  = get_device_split_info(arr);
    const int nparts        = int(device_info.size());
    const value_t* in       = arr.data().get();
    functor_t op(f);
    b200::call_workspaces ws(device_info);  // leased for this call only
    value_t* partials = ws.scratch<value_t>(size_t(nparts));
    b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
        b200::launch_reduce<value_t, functor_t>(
            in + p.offset, p.elements, identity, partials + p.ordinal, op, p, ws
        );
    });
    b200::sync_devices(b200::devices_of(device_info));
    std::vector<value_t> host(size_t(nparts), identity);
    cuda::cuda_exception::raise_if_not_success(
        cudaMemcpy(host.data(), partials, sizeof(value_t) * size_t(nparts), cudaMemcpyDeviceToHost),
        std::experimental::source_location::current(),
        "sclx::algorithm::"
    );
    cuda::set_device(caller_device);
    value_t result = identity;
    for (const value_t& partial : host) result = op(result, partial);
    return result;
}

}  // namespace sclx::algorithm
