// scalix-b200: CUDA shim.  Public surface of /root/reference scalix/cuda.hpp:41-305 and
// scalix/detail/cuda.hpp:38-116 on top of the scalix-b200 runtime (libscalix_b200.so):
//
//   * device ids are LOGICAL devices of the runtime: the physical GPUs, or N emulated
//     devices on GPU 0 when SCLX_B200_EMULATE_DEVICES=N / -DSCALIX_EMULATE_MULTIDEVICE=N
//     (the reference's compile-time switch, /root/reference CMakeLists.txt:31-44);
//   * mem_advise / mem_prefetch_async are kept for source compatibility but carry no
//     page-migration semantics: arrays are explicit per-device allocations (see array.cuh);
//   * task_scheduler::submit_task runs on ONE persistent host thread per logical device
//     (same signature as the reference's cuda_thread_pool front-end) — the launcher in
//     detail/execute_kernel.cuh uses the same threads for multi-device launch groups.
#pragma once

#include "../scalix_b200.h"
#include "shape.cuh"
#include "throw_exception.hpp"

#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <deque>
#include <fstream>
#include <functional>
#include <future>
#include <limits>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

namespace sclx::cuda {

struct cuda_exception : std::runtime_error {
    explicit cuda_exception(const std::string& msg) : std::runtime_error(msg) {}

    static void raise_if_not_success(
        cudaError_t err,
        const std::experimental::source_location& location,
        const std::string& function_prefix = "sclx::cuda::"
    ) {
        if (err != cudaSuccess) {
            throw_exception<cuda_exception>(cudaGetErrorString(err), function_prefix, location);
        }
    }
};

namespace detail {

// turns a scalix-b200 C-ABI status into the reference's exception types
inline void check_rt(
    int rc,
    const std::experimental::source_location& location
    = std::experimental::source_location::current(),
    const std::string& function_prefix = "sclx::"
) {
    if (rc == 0) return;
    const std::string msg = sclx_b200_last_error();
    if (rc == SCLX_E_INVALID_ARGUMENT)
        throw_exception<std::invalid_argument>(msg, function_prefix, location);
    if (rc < 0) throw_exception<std::runtime_error>(msg, function_prefix, location);
    throw_exception<cuda_exception>(msg, function_prefix, location);
}

inline void ensure_runtime() {
    static std::once_flag once;
    std::call_once(once, [] {
#ifdef SCALIX_EMULATE_MULTIDEVICE
        if (!std::getenv("SCLX_B200_EMULATE_DEVICES"))
            sclx_b200_rt_set_emulated_devices(SCALIX_EMULATE_MULTIDEVICE);
#endif
        check_rt(sclx_b200_rt_init());
    });
}

}  // namespace detail

// `device` is a logical device of the runtime
inline void set_device(int device) {
    detail::ensure_runtime();
    const int physical = sclx_b200_rt_physical_device(device);
    cudaError_t err    = physical < 0 ? cudaErrorInvalidDevice : cudaSetDevice(physical);
    cuda_exception::raise_if_not_success(err, std::experimental::source_location::current());
}

struct traits {
    static constexpr int cpu_device_id = cudaCpuDeviceId;

    enum mem_advise {
        set_read_mostly          = cudaMemAdviseSetReadMostly,
        unset_read_mostly        = cudaMemAdviseUnsetReadMostly,
        set_preferred_location   = cudaMemAdviseSetPreferredLocation,
        unset_preferred_location = cudaMemAdviseUnsetPreferredLocation,
        set_accessed_by          = cudaMemAdviseSetAccessedBy,
        unset_accessed_by        = cudaMemAdviseUnsetAccessedBy
    };

    struct kernel {
        static constexpr shape_t<1> default_block_shape{256};
        static constexpr size_t default_grid_size = std::numeric_limits<size_t>::max();

        static auto max_thread_count() -> const size_t& {
            static size_t max_thread_count = std::numeric_limits<size_t>::max();
            return max_thread_count;
        }
        static void set_max_thread_count(size_t count) {
            const_cast<size_t&>(max_thread_count()) = count;
        }
    };

    // number of logical devices
    static int device_count() {
        detail::ensure_runtime();
        return sclx_b200_rt_device_count();
    }

    // logical id of the calling thread's current device (emulation: always 0's physical)
    static int current_device() {
        int device = 0;
        cudaGetDevice(&device);
        detail::ensure_runtime();
        const int n = sclx_b200_rt_device_count();
        for (int d = 0; d < n; ++d)
            if (sclx_b200_rt_physical_device(d) == device) return d;
        return 0;
    }
};

// Explicit per-device allocations need no advice; kept so user code compiles.
template<class T>
void mem_advise(T*, size_t, cuda::traits::mem_advise, int) {}

// "Prefetch to the CPU" = bring the host mirror of the range up to date now;
// prefetch to a device = make sure host stores are visible to the device.
template<class T>
void mem_prefetch_async(T* begin, T* end, int device, cudaStream_t = nullptr) {
    detail::ensure_runtime();
    if (device == traits::cpu_device_id) {
        detail::check_rt(sclx_b200_rt_prefetch_to_host(
            begin, size_t(reinterpret_cast<const char*>(end) - reinterpret_cast<const char*>(begin))
        ));
    } else {
        detail::check_rt(sclx_b200_rt_before_device_access());
    }
}

inline void stream_synchronize(cudaStream_t stream = nullptr) {
    cuda_exception::raise_if_not_success(
        cudaStreamSynchronize(stream), std::experimental::source_location::current()
    );
}

inline void peak_last_error_and_throw_if_error(
    const std::experimental::source_location& location,
    const std::string& function_prefix = "sclx::cuda::"
) {
    cuda_exception::raise_if_not_success(cudaPeekAtLastError(), location, function_prefix);
}

inline void get_last_error_and_throw_if_error(
    const std::experimental::source_location& location,
    const std::string& function_prefix = "sclx::cuda::"
) {
    cuda_exception::raise_if_not_success(cudaGetLastError(), location, function_prefix);
}

struct memory_status_info {
    size_t free;
    size_t total;
};

inline memory_status_info query_memory_status(int device = 0) {
    const int before = traits::current_device();
    set_device(device);
    size_t free = 0, total = 0;
    cudaError_t err = cudaMemGetInfo(&free, &total);
    set_device(before);
    cuda_exception::raise_if_not_success(err, std::experimental::source_location::current());
    // arrays parked in the runtime's reuse pool are handed back on demand: count them as free
    free += sclx_b200_rt_pool_bytes_on(device);
    if (free > total) free = total;
    // emulated devices share one GPU
    int sharing = 0;
    for (int d = 0; d < traits::device_count(); ++d)
        sharing += sclx_b200_rt_physical_device(d) == sclx_b200_rt_physical_device(device);
    if (sharing > 1) {
        free /= size_t(sharing);
        total /= size_t(sharing);
    }
    return {free, total};
}

namespace host {
inline memory_status_info query_memory_status() {
    std::ifstream meminfo("/proc/meminfo");
    std::string key;
    size_t value = 0, total = 0, available = 0;
    std::string unit;
    while (meminfo >> key >> value) {
        std::getline(meminfo, unit);
        if (key == "MemTotal:") total = value * 1000;
        if (key == "MemAvailable:") available = value * 1000;
    }
    return {available, total};
}
}  // namespace host

// shared-ownership non-blocking stream (API of /root/reference scalix/cuda.hpp:215-268)
class stream_t {
  public:
    stream_t()                           = default;
    stream_t(const stream_t&)            = default;
    stream_t& operator=(const stream_t&) = default;
    stream_t(stream_t&&)                 = default;
    stream_t& operator=(stream_t&&)      = default;

    explicit stream_t(const cudaStream_t& stream) { *stream_ = stream; }

    static stream_t create() {
        cudaStream_t stream;
        cuda_exception::raise_if_not_success(
            cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking),
            std::experimental::source_location::current()
        );
        return stream_t{stream};
    }

    static stream_t create_for_device(int device) {
        const int current = traits::current_device();
        set_device(device);
        stream_t stream = create();
        set_device(current);
        return stream;
    }

    operator cudaStream_t() const { return *stream_; }  // NOLINT

    [[nodiscard]] const cudaStream_t& get() const { return *stream_; }

    void synchronize() const {
        cuda_exception::raise_if_not_success(
            cudaStreamSynchronize(*stream_), std::experimental::source_location::current()
        );
    }

  private:
    std::shared_ptr<cudaStream_t> stream_
        = std::shared_ptr<cudaStream_t>(new cudaStream_t{nullptr}, [](cudaStream_t* s) {
              if (*s) cudaStreamDestroy(*s);
              delete s;
          });
};

namespace detail {

// One persistent host thread per logical device, bound to that device for its whole
// life (the reference's pool spins with yield() forever,
// /root/reference scalix/detail/cuda.hpp:70-95).  A worker that has just finished a job keeps
// polling for the next one for a short while before it goes to sleep on its condition
// variable, so the launch groups of a multi-pass algorithm or of a loop of launches are handed
// over in well under a microsecond, and an idle process burns no CPU.
class device_workers {
  public:
    static device_workers& instance() {
        static device_workers pool(traits::device_count());
        return pool;
    }

    // task_scheduler::submit_task: run `f` on the device's thread, result through a future
    template<class F>
    auto submit(int device_id, F&& f) -> std::future<std::invoke_result_t<std::decay_t<F>>> {
        using result_t = std::invoke_result_t<std::decay_t<F>>;
        if (device_id < 0 || device_id >= int(workers_.size()))
            throw_exception<std::invalid_argument>("device id out of range", "sclx::cuda::");
        auto task   = std::make_shared<std::packaged_task<result_t()>>(std::forward<F>(f));
        auto future = task->get_future();
        enqueue(*workers_[device_id], [task] { (*task)(); });
        return future;
    }

    // fork-join used by launch groups: job(i) runs on the thread of devices[i] for every
    // i in [0, n); returns when all have finished and rethrows the first exception
    template<class Job>
    void run_on(const int* devices, int n, Job&& job) {
        for (int i = 0; i < n; ++i)
            if (devices[i] < 0 || devices[i] >= int(workers_.size()))
                throw_exception<std::invalid_argument>("device id out of range", "sclx::cuda::");
        struct join_t {
            std::atomic<int> remaining;
            std::mutex mutex;
            std::exception_ptr first;
        } join{{n}, {}, {}};
        for (int i = 0; i < n; ++i)
            enqueue(*workers_[devices[i]], [&join, &job, i] {
                try {
                    job(i);
                } catch (...) {
                    std::lock_guard<std::mutex> lock(join.mutex);
                    if (!join.first) join.first = std::current_exception();
                }
                join.remaining.fetch_sub(1, std::memory_order_release);
            });
        // the jobs only queue kernels (microseconds): poll, then be polite
        for (unsigned spins = 0; join.remaining.load(std::memory_order_acquire) != 0; ++spins)
            if (spins > 4096) std::this_thread::yield();
        if (join.first) std::rethrow_exception(join.first);
    }

    ~device_workers() {
        for (auto& w : workers_) {
            {
                std::lock_guard<std::mutex> lock(w->mutex);
                w->stop = true;
            }
            w->pending.fetch_add(1, std::memory_order_release);
            w->wake.notify_one();
        }
        for (auto& w : workers_) w->thread.join();
    }

  private:
    struct worker {
        std::mutex mutex;
        std::condition_variable wake;
        std::deque<std::function<void()>> queue;
        std::atomic<int> pending{0};  // queued jobs, read by the polling worker without the lock
        bool stop = false;
        std::thread thread;
    };

    static void enqueue(worker& w, std::function<void()> fn) {
        {
            std::lock_guard<std::mutex> lock(w.mutex);
            w.queue.emplace_back(std::move(fn));
        }
        w.pending.fetch_add(1, std::memory_order_release);
        w.wake.notify_one();
    }

    static long poll_microseconds() {
        if (const char* e = std::getenv("SCLX_B200_WORKER_SPIN_US")) return std::atol(e);
        // long enough to bridge one large kernel between two launch groups of a loop (a 4 GiB
        // reduce_last_dim step is 0.6 ms): the next group then finds the workers awake
        return 2000;
    }

    explicit device_workers(int count) {
        const auto poll = std::chrono::microseconds(poll_microseconds());
        for (int d = 0; d < count; ++d) {
            workers_.emplace_back(new worker);
            worker* w = workers_.back().get();
            w->thread = std::thread([w, d, poll] {
                cudaSetDevice(sclx_b200_rt_physical_device(d));
                for (;;) {
                    // hot phase: the next job usually follows within microseconds
                    const auto t0 = std::chrono::steady_clock::now();
                    while (w->pending.load(std::memory_order_acquire) == 0
                           && std::chrono::steady_clock::now() - t0 < poll) {
#if defined(__x86_64__)
                        __builtin_ia32_pause();
#endif
                    }
                    std::function<void()> job;
                    {
                        std::unique_lock<std::mutex> lock(w->mutex);
                        w->wake.wait(lock, [w] { return w->stop || !w->queue.empty(); });
                        if (w->queue.empty()) return;
                        job = std::move(w->queue.front());
                        w->queue.pop_front();
                    }
                    w->pending.fetch_sub(1, std::memory_order_relaxed);
                    job();
                }
            });
        }
    }

    std::vector<std::unique_ptr<worker>> workers_;
};

}  // namespace detail

struct task_scheduler {
    template<class F, class... Args>
    static std::future<std::invoke_result_t<F, Args...>>
    submit_task(int device_id, F&& f, Args&&... args) {
        auto bound = [f = std::forward<F>(f),
                      args_tuple = std::make_tuple(std::forward<Args>(args)...)]() mutable {
            return std::apply(f, args_tuple);
        };
        return detail::device_workers::instance().submit(device_id, std::move(bound));
    }
};

}  // namespace sclx::cuda
