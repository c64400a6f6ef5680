// scalix-b200: algorithm::count_if (surface of /root/reference
// scalix/algorithm/count_if.cuh:44-98): number of elements satisfying a predicate, as a host
// scalar.  One map_reduce_kernel per partition, partial counts summed on the host.
#pragma once
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"
#include "functional.cuh"

#include <vector>

namespace sclx::algorithm {

namespace detail {
template<class Pred>
struct predicate_to_count {
    Pred pred;
    template<class T>
    __host__ __device__ unsigned long long operator()(const T& x) const {
        return pred(x) ? 1ull : 0ull;
    }
};
struct add_counts {
    __host__ __device__ unsigned long long
    operator()(unsigned long long a, unsigned long long b) const {
        return a + b;
    }
};
}  // namespace detail

template<class T, uint Rank, class F>
__host__ size_t count_if(const array<T, Rank>& arr, F&& f) {
    using value_t = std::remove_const_t<T>;
    using count_t = unsigned long long;
    using map_t   = detail::predicate_to_count<std::decay_t<F>>;
    const int caller_device = cuda::traits::current_device();
    const auto device_info  = get_device_split_info(arr);
    const int nparts        = int(device_info.size());
    const value_t* in       = arr.data().get();
    b200::call_workspaces ws(device_info);  // leased for this call only
    count_t* partials       = ws.scratch<count_t>(size_t(nparts));
    map_t map{std::decay_t<F>(f)};
    b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
        b200::launch_map_reduce<value_t, count_t, map_t, detail::add_counts>(
            in + p.offset, p.elements, count_t(0), partials + p.ordinal, map, detail::add_counts{}, p, ws
        );
    });
    b200::sync_devices(b200::devices_of(device_info));
    std::vector<count_t> host(size_t(nparts), 0);
    cuda::cuda_exception::raise_if_not_success(
        cudaMemcpy(host.data(), partials, sizeof(count_t) * size_t(nparts), cudaMemcpyDeviceToHost),
        std::experimental::source_location::current(),
        "sclx::algorithm::"
    );
    cuda::set_device(caller_device);
    size_t total = 0;
    for (count_t c : host) total += size_t(c);
    return total;
}

}  // namespace sclx::algorithm
