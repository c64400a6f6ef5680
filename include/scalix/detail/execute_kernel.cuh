// scalix-b200: launch machinery behind kernel_handler::launch.
//
// Replaces /root/reference scalix/detail/execute_kernel.cuh:41-581.  Same extension points —
// detail::default_kernel_tag with the two `execute` member templates and
// REGISTER_SCALIX_KERNEL_TAG(Tag) stamping a distinctly named kernel pair — but:
//
//   * no stream is created or destroyed per launch: every logical device owns ONE persistent
//     non-blocking stream in the runtime;
//   * a launch group is ordered after the previous group by EVENTS on the device
//     (sclx_b200_rt_begin_launch / end_launch) and returns as soon as the kernels are queued;
//     the future returned by execute_kernel synchronises the streams that were used;
//   * launches for several devices are issued in parallel by the persistent per-device host
//     threads (cuda::detail::device_workers) instead of a freshly spawned std::async thread;
//   * grids are sized to the machine (launch_config.cuh); launch errors are checked at once
//     (the reference only notices them at stream.synchronize()).
#pragma once

#include "../array.cuh"
#include "../cuda.hpp"
#include "../kernel_info.cuh"
#include "b200/dispatch.cuh"
#include "b200/launch_config.cuh"

#include <atomic>
#include <cstdlib>
#include <iostream>
#include <mutex>
#include <string>
#include <tuple>
#include <typeinfo>
#include <vector>

// -DSCALIX_NVTX: one NVTX range per launch group, named after the kernel tag, so a timeline
// (Nsight Systems / ncu --nvtx) shows which sclx launch a kernel belongs to.  Header-only
// NVTX v3 from the CUDA toolkit; a no-op unless a profiler is attached.
#ifdef SCALIX_NVTX
#include <nvtx3/nvToolsExt.h>
#endif

namespace sclx::detail {

struct nvtx_range {
#ifdef SCALIX_NVTX
    explicit nvtx_range(const char* what, const char* tag) {
        std::string name = std::string(what) + " [" + tag + "]";
        nvtxRangePushA(name.c_str());
    }
    ~nvtx_range() { nvtxRangePop(); }
#else
    explicit nvtx_range(const char*, const char*) {}
#endif
    nvtx_range(const nvtx_range&)            = delete;
    nvtx_range& operator=(const nvtx_range&) = delete;
};

#ifdef SCALIX_DEBUG_KERNEL_LAUNCH
constexpr bool _debug_kernel_launch = true;
#else
constexpr bool _debug_kernel_launch = false;
#endif

struct _debug_kernel_mutex {
    static std::mutex& get() {
        static std::mutex m;
        return m;
    }
};

// logical devices the calling thread has queued work on since execute_kernel started
struct launch_record {
    static std::vector<int>& devices() {
        thread_local std::vector<int> used;
        return used;
    }
    static void note(int device) {
        auto& used = devices();
        for (int d : used)
            if (d == device) return;
        used.push_back(device);
    }
};

using device_info_t = std::vector<std::tuple<int, size_t, size_t>>;

// launch groups of at most this many devices are queued by the calling thread itself
// (SCLX_B200_INLINE_ISSUE overrides)
inline int inline_issue_limit() {
    static const int limit = [] {
        const char* e = std::getenv("SCLX_B200_INLINE_ISSUE");
        return e ? std::atoi(e) : 2;
    }();
    return limit;
}

// `issue(ordinal, device_id, slice_idx, slice_size, stream)` queues one kernel; called with
// the device current: inline for small groups, on the devices' persistent threads otherwise
template<class Issue>
void run_launch_group(const device_info_t& device_info, Issue&& issue) {
    cuda::detail::ensure_runtime();
    const int n = int(device_info.size());
    std::vector<int> devices(n);
    for (int i = 0; i < n; ++i) devices[i] = std::get<0>(device_info[i]);
    cuda::detail::check_rt(sclx_b200_rt_begin_launch(n, devices.data()));

    const int caller_device = cuda::traits::current_device();
    auto issue_one          = [&](int ordinal) {
        const auto& [device_id, slice_idx, slice_size] = device_info[ordinal];
        cuda::set_device(device_id);
        auto stream = static_cast<cudaStream_t>(sclx_b200_rt_stream(device_id));
        issue(ordinal, device_id, slice_idx, slice_size, stream);
        cuda::cuda_exception::raise_if_not_success(
            cudaGetLastError(), std::experimental::source_location::current(), "sclx::detail::"
        );
        cuda::detail::check_rt(sclx_b200_rt_signal_done(device_id));  // what waits poll for
    };
    if (n <= inline_issue_limit()) {
        // one or two kernels are queued faster from here than a hand-over costs
        for (int i = 0; i < n; ++i) issue_one(i);
    } else {
        // one persistent thread per device queues its kernel: N launches in parallel
        cuda::detail::device_workers::instance().run_on(devices.data(), n, issue_one);
    }
    cuda::set_device(caller_device);
    cuda::detail::check_rt(sclx_b200_rt_end_launch(n, devices.data()));
    sclx_b200_note_launches(static_cast<unsigned long long>(n));
    for (int d : devices) launch_record::note(d);
}

// which captured arrays the range kernel prefetches one trip ahead (prefetch_plan_t,
// b200/dispatch.cuh).  SCLX_B200_DISPATCH_PREFETCH=0 switches it off.
template<class Functor>
prefetch_plan_t make_prefetch_plan(const Functor& f, size_t range_elements) {
    prefetch_plan_t plan;
    static const bool enabled = [] {
        const char* e = std::getenv("SCLX_B200_DISPATCH_PREFETCH");
        return !(e && e[0] == '0');
    }();
    // small launches are latency-bound: not worth the scan
    if (!enabled || range_elements < (size_t(1) << 16)) return plan;
    void* bases[prefetch_plan_t::max_arrays];
    size_t bytes[prefetch_plan_t::max_arrays];
    const int found = sclx_b200_rt_find_captured_arrays(
        &f, sizeof(Functor) <= 1024 ? sizeof(Functor) : 1024, prefetch_plan_t::max_arrays, bases, bytes
    );
    for (int k = 0; k < found; ++k) {
        if (bytes[k] % range_elements != 0) continue;
        const size_t per = bytes[k] / range_elements;
        // 16 bytes per range element is far more often a LARGER array indexed indirectly (a
        // gather source) than an array of 16-byte elements: leave those alone
        if (per != 1 && per != 2 && per != 4 && per != 8) continue;
        plan.add(bases[k], unsigned(per));
    }
    return plan;
}

// Two launch groups in one: on every device `first` then `second`, where `second` on device i
// must see what `first` produced on other devices (all of them, or only the partitions below
// i when `lower_only`).  Issued as ONE fork-join: each device's thread queues its first kernel,
// records its event, meets the others at a host-side barrier (all events recorded), makes its
// stream wait for the events it needs and queues its second kernel.  Two plain groups would
// pay the hand-over twice and serialise ~3N event calls on the calling thread in between.
template<class First, class Second>
void run_chained_launch_groups(
    const device_info_t& device_info, bool lower_only, First&& first, Second&& second
) {
    const int n = int(device_info.size());
    bool distinct = true;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j) distinct &= std::get<0>(device_info[i]) != std::get<0>(device_info[j]);
    if (n == 1 || !distinct) {
        // one stream orders the two kernels by itself / a device listed twice would meet itself
        // at the barrier: plain groups
        run_launch_group(device_info, first);
        run_launch_group(device_info, second);
        return;
    }
    cuda::detail::ensure_runtime();
    std::vector<int> devices(n);
    for (int i = 0; i < n; ++i) devices[i] = std::get<0>(device_info[i]);
    cuda::detail::check_rt(sclx_b200_rt_begin_launch(n, devices.data()));
    const int caller_device = cuda::traits::current_device();
    auto issue = [&](auto& what, int ordinal) {
        const auto& [device_id, slice_idx, slice_size] = device_info[ordinal];
        cuda::set_device(device_id);
        what(ordinal, device_id, slice_idx, slice_size, static_cast<cudaStream_t>(sclx_b200_rt_stream(device_id)));
        cuda::cuda_exception::raise_if_not_success(
            cudaGetLastError(), std::experimental::source_location::current(), "sclx::detail::"
        );
    };
    auto wait_for_others = [&](int ordinal) {
        for (int j = 0; j < (lower_only ? ordinal : n); ++j)
            if (j != ordinal) cuda::detail::check_rt(sclx_b200_rt_wait_done(devices[ordinal], devices[j]));
    };
    if (n <= inline_issue_limit()) {
        for (int i = 0; i < n; ++i) {
            issue(first, i);
            cuda::detail::check_rt(sclx_b200_rt_record_done(devices[i]));
        }
        for (int i = 0; i < n; ++i) {
            wait_for_others(i);
            issue(second, i);
            cuda::detail::check_rt(sclx_b200_rt_signal_done(devices[i]));
        }
    } else {
        std::atomic<int> arrived{0};
        std::atomic<bool> failed{false};
        cuda::detail::device_workers::instance().run_on(devices.data(), n, [&](int i) {
            std::exception_ptr error;
            try {
                issue(first, i);
                cuda::detail::check_rt(sclx_b200_rt_record_done(devices[i]));
            } catch (...) {
                error = std::current_exception();
                failed.store(true);
            }
            // every device's event is recorded before anybody waits for one
            arrived.fetch_add(1, std::memory_order_acq_rel);
            while (arrived.load(std::memory_order_acquire) < n) {
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            }
            if (error) std::rethrow_exception(error);
            if (failed.load()) return;
            wait_for_others(i);
            issue(second, i);
            cuda::detail::check_rt(sclx_b200_rt_signal_done(devices[i]));
        });
    }
    cuda::set_device(caller_device);
    cuda::detail::check_rt(sclx_b200_rt_end_launch(n, devices.data()));
    sclx_b200_note_launches(static_cast<unsigned long long>(2 * n));
    for (int d : devices) launch_record::note(d);
}

template<class Kernel>
void allow_dynamic_shared_memory(Kernel kernel, size_t local_mem_size) {
    if (local_mem_size <= 48u * 1024u) return;
    cuda::cuda_exception::raise_if_not_success(
        cudaFuncSetAttribute(
            kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(local_mem_size)
        ),
        std::experimental::source_location::current(),
        "sclx::detail::"
    );
}

// The `execute` pair every kernel tag exposes.  `Kernels` names the __global__ templates:
//   Kernels::template range<F, RangeRank, TBRank>()          -> kernel pointer
//   Kernels::template generator<F, IndexGenerator, TBRank>() -> kernel pointer
template<class Kernels>
struct tag_executor {
    // range launch: partition p runs the sub-range whose last coordinate lies in its slice
    template<uint RangeRank, uint ThreadBlockRank, class F>
    __host__ static void execute(
        const device_info_t& device_info,
        const sclx::md_range_t<RangeRank>& global_range,
        const sclx::shape_t<ThreadBlockRank>& block_shape,
        size_t grid_size,
        size_t local_mem_size,
        F&& f
    ) {
        using functor_t = std::decay_t<F>;
        auto kernel     = Kernels::template range<functor_t, RangeRank, ThreadBlockRank>();
        const unsigned block_threads = unsigned(block_shape.elements());
        const prefetch_plan_t plan = make_prefetch_plan<functor_t>(f, global_range.elements());
        nvtx_range trace("sclx range launch", typeid(Kernels).name());
        std::unique_lock<std::mutex> debug_lock(_debug_kernel_mutex::get(), std::defer_lock);
        if constexpr (_debug_kernel_launch) {
            debug_lock.lock();
            std::cout << "Launching Scalix Kernel" << std::endl;
        }
        run_launch_group(
            device_info,
            [&](int ordinal, int device_id, size_t slice_idx, size_t slice_size, cudaStream_t stream) {
                sclx::md_range_t<RangeRank> device_range(global_range);
                device_range[RangeRank - 1] = slice_size;
                sclx::md_index_t<RangeRank> device_start_idx{};
                device_start_idx[RangeRank - 1] = slice_idx;
                const size_t elements           = device_range.elements();
                if (elements == 0) return;
                sclx::kernel_info<RangeRank, ThreadBlockRank> metadata(
                    global_range, device_range, block_shape, device_start_idx, ordinal
                );
                if constexpr (_debug_kernel_launch) {
                    std::cout << "  Launching kernel on device " << device_id << std::endl;
                    std::cout << "      Device range: " << device_range << std::endl;
                    std::cout << "      Device start index: " << device_start_idx << std::endl;
                    std::cout << "      Global range: " << global_range << std::endl;
                }
                (void)device_id;
                unsigned grid = 1;
                cuda::cuda_exception::raise_if_not_success(
                    sclx::b200::range_launch_grid(
                        kernel, elements, block_threads, local_mem_size, grid_size, &grid,
                        sclx::cuda::traits::kernel::max_thread_count(),
                        global_range[RangeRank - 1] ? global_range.elements() / global_range[RangeRank - 1] : 1
                    ),
                    std::experimental::source_location::current(),
                    "sclx::detail::"
                );
                allow_dynamic_shared_memory(kernel, local_mem_size);
                kernel<<<grid, block_threads, local_mem_size, stream>>>(f, metadata, plan);
            }
        );
        if constexpr (_debug_kernel_launch)
            std::cout << "Kernel launch complete" << std::endl << std::endl;
    }

    // index-generator launch: every partition walks the WHOLE generator range and keeps the
    // indices whose last coordinate it owns (replicated execution with owner filter)
    template<class IndexGenerator, uint ThreadBlockRank, class F>
    __host__ static void execute(
        const device_info_t& device_info,
        IndexGenerator&& index_generator,
        const sclx::shape_t<ThreadBlockRank>& block_shape,
        size_t grid_size,
        size_t local_mem_size,
        F&& f
    ) {
        using functor_t   = std::decay_t<F>;
        using generator_t = std::decay_t<IndexGenerator>;
        constexpr uint range_rank = generator_t::range_rank;
        auto kernel = Kernels::template generator<functor_t, generator_t, ThreadBlockRank>();
        const unsigned block_threads = unsigned(block_shape.elements());
        nvtx_range trace("sclx index-generator launch", typeid(Kernels).name());
        std::unique_lock<std::mutex> debug_lock(_debug_kernel_mutex::get(), std::defer_lock);
        if constexpr (_debug_kernel_launch) {
            debug_lock.lock();
            std::cout << "Launching Scalix Kernel using IndexGenerator" << std::endl;
        }
        const auto generator_range = index_generator.range();
        run_launch_group(
            device_info,
            [&](int ordinal, int device_id, size_t slice_idx, size_t slice_size, cudaStream_t stream) {
                const size_t elements = generator_range.elements();
                if (elements == 0 || slice_size == 0) return;
                sclx::kernel_info<range_rank, ThreadBlockRank> metadata(
                    generator_range, generator_range, block_shape, {}, ordinal
                );
                if constexpr (_debug_kernel_launch) {
                    std::cout << "  Launching kernel on device " << device_id << std::endl;
                    std::cout << "      Device slice size: " << slice_size << std::endl;
                    std::cout << "      Device slice start index: " << slice_idx << std::endl;
                }
                (void)device_id;
                unsigned grid = 1;
                cuda::cuda_exception::raise_if_not_success(
                    sclx::b200::range_launch_grid(
                        kernel, elements, block_threads, local_mem_size, grid_size, &grid,
                        sclx::cuda::traits::kernel::max_thread_count()
                    ),
                    std::experimental::source_location::current(),
                    "sclx::detail::"
                );
                allow_dynamic_shared_memory(kernel, local_mem_size);
                kernel<<<grid, block_threads, local_mem_size, stream>>>(
                    f, metadata, index_generator, slice_idx, slice_size
                );
            }
        );
        if constexpr (_debug_kernel_launch)
            std::cout << "Kernel launch complete" << std::endl << std::endl;
    }
};

struct default_kernels {
    template<class F, uint RangeRank, uint ThreadBlockRank>
    static auto range() {
        return &sclx_kernel<F, RangeRank, ThreadBlockRank>;
    }
    template<class F, class IndexGenerator, uint ThreadBlockRank>
    static auto generator() {
        return &sclx_kernel<F, IndexGenerator, ThreadBlockRank>;
    }
};

struct default_kernel_tag : tag_executor<default_kernels> {};

}  // namespace sclx::detail

// Stamps `sclx_<Tag>` kernels (so profilers show a distinct name per tag) and a struct `Tag`
// with the default_kernel_tag interface; use as `handler.launch<Tag>(...)`.
#define REGISTER_SCALIX_KERNEL_TAG(Tag)                                                          \
    template<class F, uint RangeRank, uint ThreadBlockRank>                                      \
    __global__ void sclx_##Tag(                                                                  \
        F f,                                                                                     \
        sclx::kernel_info<RangeRank, ThreadBlockRank> metadata,                                  \
        sclx::detail::prefetch_plan_t plan                                                       \
    ) {                                                                                          \
        sclx::detail::range_kernel_body(f, metadata, plan);                                      \
    }                                                                                            \
    template<class F, class IndexGenerator, uint ThreadBlockRank>                                \
    __global__ void sclx_##Tag(                                                                  \
        F f,                                                                                     \
        sclx::kernel_info<IndexGenerator::range_rank, ThreadBlockRank> metadata,                 \
        IndexGenerator idx_gen,                                                                  \
        sclx::index_t slice_idx,                                                                 \
        sclx::index_t slice_size                                                                 \
    ) {                                                                                          \
        sclx::detail::generator_kernel_body(f, metadata, idx_gen, slice_idx, slice_size);        \
    }                                                                                            \
    struct Tag##_sclx_kernels {                                                                  \
        template<class F, uint RangeRank, uint ThreadBlockRank>                                  \
        static auto range() {                                                                    \
            return &sclx_##Tag<F, RangeRank, ThreadBlockRank>;                                   \
        }                                                                                        \
        template<class F, class IndexGenerator, uint ThreadBlockRank>                            \
        static auto generator() {                                                                \
            return &sclx_##Tag<F, IndexGenerator, ThreadBlockRank>;                              \
        }                                                                                        \
    };                                                                                           \
    struct Tag : sclx::detail::tag_executor<Tag##_sclx_kernels> {}
