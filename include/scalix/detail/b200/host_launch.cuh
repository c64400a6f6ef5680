// scalix-b200: host-side launchers of the framework-owned kernels on partitioned arrays.
// Used by fill.cuh, iota.cuh, assign_array.cuh and algorithm/*.cuh: each partition of an
// array gets one launch on its owner's persistent stream, grouped and ordered exactly like a
// kernel_handler::launch (detail::run_launch_group).
#pragma once

#include "../../array.cuh"
#include "../execute_kernel.cuh"
#include "reduce.cuh"
#include "scan_stream.cuh"
#include "stream_kernels.cuh"

#include <exception>
#include <type_traits>
#include <vector>

namespace sclx::b200 {

struct partition_t {
    int ordinal;      // position in the array's split
    int device;       // logical device
    size_t offset;    // first element (relative to the array)
    size_t elements;  // element count
    cudaStream_t stream;
};

// fn(partition_t) with the partition's device current; one launch group
template<class T, uint Rank, class Fn>
void for_each_partition(
    const array<T, Rank>& arr,
    const sclx::detail::device_info_t& device_info,
    Fn&& fn
) {
    const size_t last = arr.shape()[Rank - 1];
    const size_t lead = last ? arr.elements() / last : 0;
    sclx::detail::run_launch_group(
        device_info,
        [&](int ordinal, int device_id, size_t slice_idx, size_t slice_size, cudaStream_t stream) {
            fn(partition_t{ordinal, device_id, slice_idx * lead, slice_size * lead, stream});
        }
    );
}

// two dependent kernels per partition in one fork-join (sclx::detail::run_chained_launch_groups)
template<class T, uint Rank, class First, class Second>
void for_each_partition_chained(
    const array<T, Rank>& arr,
    const sclx::detail::device_info_t& device_info,
    bool lower_only,
    First&& first,
    Second&& second
) {
    const size_t last = arr.shape()[Rank - 1];
    const size_t lead = last ? arr.elements() / last : 0;
    auto wrap = [lead](auto& fn) {
        return [&fn, lead](int ordinal, int device_id, size_t slice_idx, size_t slice_size, cudaStream_t stream) {
            fn(partition_t{ordinal, device_id, slice_idx * lead, slice_size * lead, stream});
        };
    };
    sclx::detail::run_chained_launch_groups(device_info, lower_only, wrap(first), wrap(second));
}

template<class T, uint Rank, class Fn>
void for_each_partition(const array<T, Rank>& arr, Fn&& fn) {
    for_each_partition(arr, get_device_split_info(arr), std::forward<Fn>(fn));
}

inline std::vector<int> devices_of(const sclx::detail::device_info_t& info) {
    std::vector<int> d;
    for (const auto& t : info) d.push_back(std::get<0>(t));
    return d;
}

inline void sync_devices(const std::vector<int>& devices) {
    cuda::detail::check_rt(sclx_b200_rt_sync_devices(int(devices.size()), devices.data()));
}

// Workspace leases of ONE algorithm call: a kernel workspace (partials, tickets, look-back
// descriptors) per partition's device and any number of small peer-visible scratch blocks.
// Everything goes back to the runtime's free lists when the call ends, so two calls in
// flight on different host threads never share a buffer — the reference keeps these
// function-local (/root/reference scalix/algorithm/reduce_last_dim.cuh:180-211,
// scalix/algorithm/inclusive_scan.cuh:75-123).
class call_workspaces {
  public:
    explicit call_workspaces(const sclx::detail::device_info_t& device_info) {
        leases_.reserve(device_info.size());
        try {
            for (const auto& part : device_info) {
                void* lease = nullptr;
                cuda::detail::check_rt(sclx_b200_rt_workspace_acquire(std::get<0>(part), &lease));
                leases_.push_back(lease);
            }
        } catch (...) {
            release();
            throw;
        }
    }
    call_workspaces(const call_workspaces&)            = delete;
    call_workspaces& operator=(const call_workspaces&) = delete;
    ~call_workspaces() { release(); }

    // the lease of partition `ordinal`
    void* lease(int ordinal) const { return leases_[size_t(ordinal)]; }

    // `count` elements of peer-visible scratch (per-partition totals / partial rows)
    template<class T>
    T* scratch(size_t count) {
        void* ptr = nullptr;
        cuda::detail::check_rt(sclx_b200_rt_scratch_acquire(count * sizeof(T), &ptr));
        scratch_.push_back(ptr);
        return static_cast<T*>(ptr);
    }

  private:
    void release() {
        // a call that ends in an exception may leave tickets / arrival counters half-way: the
        // runtime re-zeroes such blocks before it hands them out again
        const int poisoned = std::uncaught_exceptions() > 0 ? 1 : 0;
        for (void* lease : leases_) sclx_b200_rt_workspace_release(lease, poisoned);
        for (void* ptr : scratch_) sclx_b200_rt_scratch_release(ptr, poisoned);
        leases_.clear();
        scratch_.clear();
    }
    std::vector<void*> leases_;
    std::vector<void*> scratch_;
};

inline int sm_count_of(int logical_device) {
    const int sms = sclx_b200_rt_sm_count(logical_device);
    return sms > 0 ? sms : kSmCount;
}

// ---- fill / iota ------------------------------------------------------------------------
template<class T>
void launch_fill(T* out, size_t n, const T& value, const partition_t& p) {
    if (n == 0) return;
    constexpr int V   = vec_traits<T>::width;
    const size_t work = (n / V + 1023) / 1024;  // one CTA per 256 x 4 vectors, up to kStreamWaves waves
    fill_kernel<T, 256, 4>
        <<<persistent_grid(work, 8, sm_count_of(p.device), kStreamWaves), 256, 0, p.stream>>>(out, n, value);
}

template<class T>
void launch_iota(T* out, size_t n, T start, size_t first_index, const partition_t& p) {
    if (n == 0) return;
    constexpr int V   = vec_traits<T>::width;
    const size_t work = (n / V + 1023) / 1024;
    iota_kernel<T, 256, 4><<<persistent_grid(work, 8, sm_count_of(p.device), kStreamWaves), 256, 0, p.stream>>>(
        out, n, start, first_index
    );
}

// ---- whole-partition reduce (op must be associative and commutative) ------------------------
template<class T, class Op>
void launch_reduce(
    const T* in, size_t n, T identity, T* out, Op op, const partition_t& p, const call_workspaces& ws_pool
) {
    constexpr int THREADS = 256, VECS = 8;
    constexpr int V       = vec_traits<T>::width;
    const size_t work     = (n / V + THREADS - 1) / THREADS;
    const unsigned grid   = persistent_grid(work, 8, sm_count_of(p.device));
    void *partials = nullptr, *ticket = nullptr;
    cuda::detail::check_rt(sclx_b200_rt_workspace_reduce(
        ws_pool.lease(p.ordinal), size_t(grid) * sizeof(T), &partials, &ticket
    ));
    reduce_workspace_t ws{partials, static_cast<unsigned*>(ticket)};
    if (is_aligned16(in))
        reduce_kernel<T, Op, THREADS, VECS, true>
            <<<grid, THREADS, 0, p.stream>>>(in, n, identity, out, ws, op);
    else
        reduce_kernel<T, Op, THREADS, VECS, false>
            <<<grid, THREADS, 0, p.stream>>>(in, n, identity, out, ws, op);
}

// ---- reduce along the last dimension for one partition ------------------------------------
template<class T, class Op>
void launch_reduce_last_dim(
    const T* in, size_t rows, size_t cols, T identity, T* out, Op op, const partition_t& p,
    const call_workspaces& ws_pool, const row_exchange_t& xchg = row_exchange_t{}
) {
    // geometry: profiles/r01_rld_tune.txt (16 loads in flight, 2 accumulator sets, 4 CTAs/SM)
    constexpr int THREADS = 256, UNROLL = 16, ACCS = 2;
    constexpr int V       = vec_traits<T>::width;
    if (rows == 0) return;
    const bool vec_ok  = (rows % V == 0) && is_aligned16(in);
    const size_t vw    = vec_ok ? V : 1;
    const size_t slots = rows / vw;
    const size_t groups = slots <= size_t(THREADS) ? THREADS / slots : 1;
    unsigned grid_y = slots <= size_t(THREADS) ? 1u : unsigned((slots + THREADS - 1) / THREADS);
    if (grid_y > 65535u)
        throw_exception<std::invalid_argument>(
            "leading extent too large for reduce_last_dim", "sclx::algorithm::"
        );
    const size_t col_work = (cols + groups - 1) / groups;
    const int sms         = sm_count_of(p.device);
    unsigned grid_x       = persistent_grid(col_work, 4, sms);
    if (grid_y > 1) {
        unsigned per_band = (unsigned(sms) * 4 + grid_y - 1) / grid_y;
        if (per_band < 1) per_band = 1;
        if (grid_x > per_band) grid_x = per_band;
    }
    void *partials = nullptr, *ticket = nullptr;
    cuda::detail::check_rt(sclx_b200_rt_workspace_reduce(
        ws_pool.lease(p.ordinal), size_t(grid_x) * rows * sizeof(T), &partials, &ticket
    ));
    reduce_workspace_t ws{partials, static_cast<unsigned*>(ticket)};
    const size_t smem = size_t(THREADS) * vw * sizeof(T);
    dim3 grid(grid_x, grid_y);
    if constexpr (V > 1) {
        if (vec_ok) {
            reduce_last_dim_kernel<T, Op, THREADS, UNROLL, V, ACCS>
                <<<grid, THREADS, smem, p.stream>>>(in, rows, cols, identity, out, ws, op, xchg);
            return;
        }
    }
    reduce_last_dim_kernel<T, Op, THREADS, UNROLL, 1, ACCS>
        <<<grid, THREADS, smem, p.stream>>>(in, rows, cols, identity, out, ws, op, xchg);
}

// ---- inclusive scan of one partition ---------------------------------------------------------
// geometry per element width: scan_stream_geometry (scan_stream.cuh)
template<class T, class Op>
void launch_inclusive_scan(
    const T* in, T* out, size_t n, const T* lower_totals, int n_lower, T* total_out, Op op,
    const partition_t& p, const call_workspaces& ws_pool
) {
    static_assert(
        std::is_trivially_copyable_v<T> && (sizeof(T) <= 4 || sizeof(T) == 8) && (16 % sizeof(T) == 0),
        "inclusive_scan kernels support trivially copyable element types of 1, 2, 4 or 8 bytes"
    );
    if (n == 0) return;
    using geo  = scan_stream_geometry<T>;
    using scfg = typename geo::config;
    using rcfg    = scan_config<T, 256, 6>;
    using desc_io = tile_desc<T>;

    const bool aligned      = is_aligned16(in) && is_aligned16(out);
    const size_t tile_items = aligned ? scfg::tile_items : rcfg::tile_items;
    const size_t tiles      = (n + tile_items - 1) / tile_items;
    if (tiles > 0x7fffffffull)
        throw_exception<std::invalid_argument>("partition too large", "sclx::algorithm::");
    void *desc_raw = nullptr, *counters = nullptr;
    unsigned epoch = 0;
    cuda::detail::check_rt(sclx_b200_rt_workspace_scan(
        ws_pool.lease(p.ordinal), tiles, int(desc_io::bytes), &desc_raw, &epoch, &counters
    ));
    auto* desc = static_cast<typename desc_io::word_t*>(desc_raw);
    scan_carry_t carry{lower_totals, n_lower};
    if (aligned) {
        auto kernel
            = inclusive_scan_stream_kernel<T, Op, geo::vecs, geo::chunks, geo::stages, geo::w_red,
                                           geo::w_scan, geo::lb_k>;
        cuda::cuda_exception::raise_if_not_success(
            cudaFuncSetAttribute(
                kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(scfg::smem_bytes)
            ),
            std::experimental::source_location::current(),
            "sclx::algorithm::"
        );
        const size_t sms    = size_t(sm_count_of(p.device));
        const unsigned grid = unsigned(tiles < sms ? tiles : sms);
        kernel<<<grid, scfg::threads, scfg::smem_bytes, p.stream>>>(
            in, out, n, desc, epoch, carry, total_out,
            static_cast<scan_stream_counters*>(counters), op
        );
    } else {
        inclusive_scan_kernel<T, Op, 256, 6, false><<<unsigned(tiles), 256, 0, p.stream>>>(
            in, out, n, desc, epoch, carry, total_out, op
        );
    }
}

// ---- map-reduce of one partition (count_if, extrema) ------------------------------------------
template<class T, class R, class Map, class Op>
void launch_map_reduce(
    const T* in, size_t n, R identity, R* out, Map map, Op op, const partition_t& p,
    const call_workspaces& ws_pool
) {
    constexpr int THREADS = 256, UNROLL = 8;
    const size_t work   = (n + THREADS - 1) / THREADS;
    const unsigned grid = persistent_grid(work, 8, sm_count_of(p.device));
    void *partials = nullptr, *ticket = nullptr;
    cuda::detail::check_rt(sclx_b200_rt_workspace_reduce(
        ws_pool.lease(p.ordinal), size_t(grid) * sizeof(R), &partials, &ticket
    ));
    reduce_workspace_t ws{partials, static_cast<unsigned*>(ticket)};
    map_reduce_kernel<T, R, Map, Op, THREADS, UNROLL>
        <<<grid, THREADS, 0, p.stream>>>(in, n, identity, out, ws, map, op);
}

// ---- out[i] = gen(first + i) over one partition ------------------------------------------------
template<class R, class Gen>
void launch_generate(R* out, size_t n, size_t first, Gen gen, const partition_t& p) {
    if (n == 0) return;
    constexpr int THREADS = 256, UNROLL = 8;
    const size_t work = (n + size_t(THREADS) * UNROLL - 1) / (size_t(THREADS) * UNROLL);
    generate_kernel<R, Gen, THREADS, UNROLL>
        <<<persistent_grid(work, 8, sm_count_of(p.device), kStreamWaves), THREADS, 0, p.stream>>>(
            out, n, first, gen
        );
}

}  // namespace sclx::b200
