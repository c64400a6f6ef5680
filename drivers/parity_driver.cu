// parity_driver — one source, two builds:
//   drivers/bin/parity_driver_ours   compiled against include/            (this repository)
//   oracle/_ref/parity_driver_ref    compiled against /root/reference     (the UNMODIFIED
//                                    reference headers, Unified-Memory build; oracle/Makefile)
// It uses ONLY the public sclx:: API, so both builds run the same program.  tests/ run both on
// the GPU box and compare the dumped outputs (bit-exact for integer work, 1e-5 relative for
// float sums); bench.py --impl reference times the reference build with the bench_* commands.
//
//   parity_driver dump  <case> <n> <outfile>          run a parity case, dump raw result bytes
//   parity_driver bench <case> <n> <steps> <warmup>   print {"ms_per_step": ..} (wall clock
//                                                      around the blocking public API call)
//   parity_driver bench_e2e rld_f32 <n> <steps> <warmup>
//                                                      same call, but every step starts with the
//                                                      input HOST-resident and ends with the result
//                                                      read on the host (copies inside the clock)
// cases: scan_i32 scan_i64 rld_f32 rld_u32max axpy_f32 axpy_f32_r3 gather_f32 fill_iota serialize
//        gather_x_f32 gather_x_f64 gather_x_skew errors algos concurrent reshape_flatten prefetch
//        copy_range tuple_launch task_scheduler future gen_hist scaling_perf serialize_file
#include <scalix/algorithm/count_if.cuh>
#include <scalix/algorithm/elementwise_reduce.cuh>
#include <scalix/algorithm/extrema.cuh>
#include <scalix/algorithm/inclusive_scan.cuh>
#include <scalix/algorithm/reduce.cuh>
#include <scalix/algorithm/reduce_last_dim.cuh>
#include <scalix/algorithm/transform.cuh>
#include <scalix/assign_array.cuh>
#include <scalix/fill.cuh>
#include <scalix/iota.cuh>
#include <scalix/scalix.cuh>

#if defined(__has_include)
#if __has_include(<cereal/archives/binary.hpp>)
#include <cereal/archives/binary.hpp>
#include <sstream>
#define DRIVER_HAS_CEREAL 1
#endif
#endif

#if defined(__has_include)
#if __has_include(<scalix/algorithm/gather.cuh>)
#include <scalix/algorithm/gather.cuh>  // extension of this repository; the reference build uses the lambda idiom
#define DRIVER_HAS_GATHER 1
#endif
#endif

#include <thrust/execution_policy.h>
#include <thrust/reduce.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <unistd.h>

// same generator as oracle/scalix_oracle.c (oracle_splitmix64): value for (seed, index)
__host__ __device__ inline unsigned long long splitmix64(unsigned long long seed, unsigned long long i) {
    unsigned long long z = seed + (i + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ inline float uniform01(unsigned long long seed, unsigned long long i) {
    return float(splitmix64(seed, i) >> 40) * (1.0f / 16777216.0f);
}
__host__ __device__ inline float symmetric(unsigned long long seed, unsigned long long i) {
    return uniform01(seed, i) * 2.0f - 1.0f;
}

struct is_even {
    __host__ __device__ bool operator()(const int& v) const { return (v & 1) == 0; }
};

template<class T, uint Rank, class Gen>
void generate(sclx::array<T, Rank>& arr, Gen gen) {
    const auto shape = arr.shape();
    sclx::execute_kernel([&](sclx::kernel_handler& handler) {
        handler.launch(
            sclx::md_range_t<Rank>(shape),
            arr,
            [=] __device__(const sclx::md_index_t<Rank>& idx, const auto&) {
                arr[idx] = gen(idx.as_linear(shape));
            }
        );
    }).get();
}

template<class T, uint Rank>
int dump(const sclx::array<T, Rank>& arr, const char* path) {
    std::vector<T> host(arr.elements());
    cudaDeviceSynchronize();
    if (cudaMemcpy(host.data(), arr.data().get(), sizeof(T) * host.size(), cudaMemcpyDefault) != cudaSuccess) {
        std::fprintf(stderr, "dump: cudaMemcpy failed\n");
        return 3;
    }
    FILE* f = std::fopen(path, "wb");
    if (!f) return 4;
    std::fwrite(host.data(), sizeof(T), host.size(), f);
    std::fclose(f);
    return 0;
}

struct timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double ms() const {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
};

// Device-side clock of the step for OUR build: one CUDA event pair per logical device on its
// persistent stream (the runtime's C ABI hands the stream out); device time of a step = the
// longest span over the devices.  The reference build creates and destroys its streams inside
// every call, so there is nothing to record on: it reports the wall clock only.
struct device_clock {
#ifdef DRIVER_HAS_GATHER
    std::vector<cudaEvent_t> start, stop;
    std::vector<cudaStream_t> streams;
    std::vector<int> physical;
    double total_ms = 0;
    device_clock() {
        const int n = sclx_b200_rt_device_count();
        for (int d = 0; d < n; ++d) {
            physical.push_back(sclx_b200_rt_physical_device(d));
            streams.push_back(static_cast<cudaStream_t>(sclx_b200_rt_stream(d)));
            cudaSetDevice(physical.back());
            cudaEvent_t a, b;
            cudaEventCreate(&a);
            cudaEventCreate(&b);
            start.push_back(a);
            stop.push_back(b);
        }
        cudaSetDevice(physical.empty() ? 0 : physical[0]);
    }
    void before() {
        for (size_t d = 0; d < streams.size(); ++d) cudaEventRecord(start[d], streams[d]);
    }
    void after() {
        float longest = 0;
        for (size_t d = 0; d < streams.size(); ++d) cudaEventRecord(stop[d], streams[d]);
        for (size_t d = 0; d < streams.size(); ++d) {
            cudaEventSynchronize(stop[d]);
            float ms = 0;
            if (cudaEventElapsedTime(&ms, start[d], stop[d]) == cudaSuccess && ms > longest) longest = ms;
        }
        total_ms += longest;
    }
    static unsigned long long launches() { return sclx_b200_launch_count(); }
    static constexpr bool available = true;
#else
    double total_ms = 0;
    void before() {}
    void after() {}
    static unsigned long long launches() { return 0; }
    static constexpr bool available = false;
#endif
};

template<class Step>
int bench(Step&& step, int steps, int warmup) {
    for (int i = 0; i < warmup; ++i) step();
    cudaDeviceSynchronize();
    const unsigned long long l0 = device_clock::launches();
    timer t;
    for (int i = 0; i < steps; ++i) step();
    const double wall = t.ms();
    const unsigned long long launches = device_clock::launches() - l0;
    // second pass with the per-device event pairs (not part of the wall clock above)
    device_clock clock;
    for (int i = 0; i < steps && device_clock::available; ++i) {
        clock.before();
        step();
        clock.after();
    }
    cudaDeviceSynchronize();
    std::printf(
        "{\"ms_per_step\": %.6f, \"steps\": %d, \"warmup\": %d, \"device_ms_per_step\": %s, \"launches\": %llu, "
        "\"devices\": %d}\n",
        wall / steps, steps, warmup,
        device_clock::available ? std::to_string(clock.total_ms / steps).c_str() : "null", launches,
        sclx::cuda::traits::device_count()
    );
    return 0;
}

// a generator for the index-generator launch and the random-read kernels: range index ->
// pseudo-random index of `target` (same interface as the reference's example generator,
// /root/reference examples/utilities/random_index_generator.cuh:78-116, but stateless)
template<uint RangeRank, uint IndexRank>
class hash_index_generator {
  public:
    static constexpr uint range_rank = RangeRank;
    static constexpr uint index_rank = IndexRank;
    __host__ __device__ hash_index_generator(sclx::shape_t<RangeRank> range, sclx::shape_t<IndexRank> target, unsigned long long seed)
        : range_(range), target_(target), seed_(seed) {}
    __host__ __device__ sclx::md_index_t<IndexRank> operator()(const sclx::md_index_t<RangeRank>& index) const {
        const size_t lin = index.as_linear(range_);
        return sclx::md_index_t<IndexRank>::create_from_linear(splitmix64(seed_, lin) % target_.elements(), target_);
    }
    __host__ __device__ const sclx::md_range_t<RangeRank>& range() const { return range_; }
    __host__ __device__ const sclx::md_range_t<IndexRank>& index_range() const { return target_; }

  private:
    sclx::md_range_t<RangeRank> range_;
    sclx::md_range_t<IndexRank> target_;
    unsigned long long seed_;
};

int write_file(const char* path, const void* data, size_t bytes) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return 4;
    std::fwrite(data, 1, bytes, f);
    std::fclose(f);
    return 0;
}

// bench_e2e: every step starts with the [64, n] input resident on the HOST (the array was
// prefetched to the CPU and overwritten there from a staging buffer, outside the clock) and
// ends when the 64 results have been read on the host.  Inside the clock: whatever the
// implementation needs to get the data to the GPUs (Unified-Memory migration in the reference
// build, the mirror's write-back here), the reduction, and the read-back.
int bench_e2e_rld(size_t n, int steps, int warmup) {
    sclx::array<float, 2> a({64, n});
    sclx::array<float, 1> r({64});
    const size_t elems = a.elements();
    float* staging = nullptr;
    if (cudaMallocHost(&staging, elems * sizeof(float)) != cudaSuccess) return 3;
    {
        // same values as the rld_f32 case, generated on the device once
        generate(a, [] __device__(size_t i) { return uniform01(0x5CA11, i); });
        cudaDeviceSynchronize();
        if (cudaMemcpy(staging, a.data().get(), elems * sizeof(float), cudaMemcpyDefault) != cudaSuccess) return 3;
    }
    double total_ms = 0, checksum = 0;
    for (int i = 0; i < warmup + steps; ++i) {
        a.prefetch_async({sclx::cuda::traits::cpu_device_id}).get();
        std::memcpy(a.data().get(), staging, elems * sizeof(float));  // host-resident input
        timer t;
        sclx::algorithm::reduce_last_dim(a, r, 0.f, sclx::algorithm::plus<>());
        double s = 0;
        for (size_t k = 0; k < 64; ++k) s += r[k];
        const double ms = t.ms();
        if (i >= warmup) total_ms += ms;
        checksum = s;
    }
    cudaFreeHost(staging);
    std::printf(
        "{\"ms_per_step\": %.6f, \"steps\": %d, \"warmup\": %d, \"h2d_bytes_per_step\": %zu, "
        "\"d2h_bytes_per_step\": %zu, \"checksum\": %.9g, \"devices\": %d}\n",
        total_ms / steps, steps, warmup, elems * sizeof(float), size_t(64) * sizeof(float), checksum,
        sclx::cuda::traits::device_count()
    );
    return 0;
}


// ---- cases --------------------------------------------------------------------------------
// every case: build inputs with the public API, run the public API call under test; `out`
// receives the array to dump; `step` re-runs only the call under test (for bench)

template<class T>
int gather_x(bool do_dump, size_t n, const char* path, int steps, int warmup, bool skew = false) {
    sclx::array<T, 1> src({4 * n}), out({n});
    sclx::array<size_t, 1> index({n});
    // skew: every index falls into the first eighth of the source (one owner's queue overflows)
    const size_t m = skew ? n / 2 : 4 * n;
    generate(src, [] __device__(size_t i) { return T(uniform01(0xABC, i)) + T(i >> 20); });
    generate(index, [=] __device__(size_t i) { return size_t(splitmix64(0x6A7, i) % m); });
    auto step = [&] {
#ifdef DRIVER_HAS_GATHER
        auto mode         = sclx::algorithm::gather_mode::automatic;
        const char* force = std::getenv("SCLX_GATHER_MODE");
        if (force && !std::strcmp(force, "exchange")) mode = sclx::algorithm::gather_mode::exchange;
        if (force && !std::strcmp(force, "direct")) mode = sclx::algorithm::gather_mode::direct;
        sclx::algorithm::gather(src, index, out, mode);
#else
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{out.shape()},
                out,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) { out[idx] = src[index[idx]]; }
            );
        }).get();
#endif
    };
    step();
    return do_dump ? dump(out, path) : bench(step, steps, warmup);
}

int run_case(const std::string& what, const std::string& name, size_t n, const char* path, int steps, int warmup) {
    const bool do_dump = what == "dump";
    if (name == "scan_i32") {
        // BASELINE.json configs[0] with the wrap-around start of BASELINE.md
        sclx::array<int, 1> a({n});
        sclx::iota(a, 0x7FFF0000);
        sclx::array<int, 1> r({n});
        auto step = [&] { sclx::algorithm::inclusive_scan(a, r, 0, sclx::algorithm::plus<>()); };
        step();
        return do_dump ? dump(r, path) : bench(step, steps, warmup);
    }
    if (name == "scan_i64") {
        // configs[4] data: SplitMix64 & 0xFFFF, rank-3 storage {2, 2, n/4} scanned flattened
        sclx::array<long long, 3> a({2, 2, n / 4});
        generate(a, [] __device__(size_t i) { return (long long)(splitmix64(0x5CA5, i) & 0xFFFF); });
        sclx::array<long long, 3> r({2, 2, n / 4});
        auto step = [&] { sclx::algorithm::inclusive_scan(a, r, 0ll, sclx::algorithm::plus<>()); };
        step();
        return do_dump ? dump(r, path) : bench(step, steps, warmup);
    }
    if (name == "rld_f32") {
        // configs[1]: float32 [64, n] sum along the last dimension
        sclx::array<float, 2> a({64, n});
        generate(a, [] __device__(size_t i) { return uniform01(0x5CA11, i); });
        sclx::array<float, 1> r({64});
        auto step = [&] { sclx::algorithm::reduce_last_dim(a, r, 0.f, sclx::algorithm::plus<>()); };
        step();
        return do_dump ? dump(r, path) : bench(step, steps, warmup);
    }
    if (name == "rld_u32max") {
        // the reference example's check (examples/reduce_last_dim.cu:55-84) on {3,3,3,n}
        sclx::array<unsigned, 4> a({3, 3, 3, n});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<4>{a.shape()},
                a,
                [=] __device__(const sclx::md_index_t<4>& index, const auto&) {
                    a[index] = unsigned(splitmix64(index[3], 7) % 1000003u) + unsigned(index[0]);
                }
            );
        }).get();
        sclx::array<unsigned, 3> r({3, 3, 3});
        auto step = [&] { sclx::algorithm::reduce_last_dim(a, r, 0u, sclx::algorithm::max<>()); };
        step();
        return do_dump ? dump(r, path) : bench(step, steps, warmup);
    }
    if (name == "axpy_f32") {
        // configs[2]: y = a*x + y through a user lambda on execute_kernel
        sclx::array<float, 1> x({n}), y({n});
        generate(x, [] __device__(size_t i) { return symmetric(1, i); });
        generate(y, [] __device__(size_t i) { return symmetric(2, i); });
        const float alpha = 1.5f;
        auto step = [&] {
            sclx::execute_kernel([&](sclx::kernel_handler& handler) {
                handler.launch(
                    sclx::md_range_t<1>{y.shape()},
                    y,
                    [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                        y[idx] = alpha * x[idx] + y[idx];
                    }
                );
            }).get();
        };
        step();
        return do_dump ? dump(y, path) : bench(step, steps, warmup);
    }
    if (name == "axpy_f32_r3") {
        sclx::array<float, 3> x({4, 4, n / 16}), y({4, 4, n / 16});
        generate(x, [] __device__(size_t i) { return symmetric(3, i); });
        generate(y, [] __device__(size_t i) { return symmetric(4, i); });
        const float alpha = -0.75f;
        auto step = [&] {
            sclx::execute_kernel([&](sclx::kernel_handler& handler) {
                handler.launch(
                    sclx::md_range_t<3>{y.shape()},
                    y,
                    [=] __device__(const sclx::md_index_t<3>& idx, const auto&) {
                        y[idx] = alpha * x[idx] + y[idx];
                    }
                );
            }).get();
        };
        step();
        return do_dump ? dump(y, path) : bench(step, steps, warmup);
    }
    if (name == "gather_f32") {
        // configs[3]: out[i] = src[index[i]], n indices into a 4n-element source
        sclx::array<float, 1> src({4 * n}), out({n});
        sclx::array<size_t, 1> index({n});
        const size_t m = 4 * n;
        generate(src, [] __device__(size_t i) { return uniform01(0xABC, i); });
        generate(index, [=] __device__(size_t i) { return size_t(splitmix64(0x6A7, i) % m); });
        auto step = [&] {
            sclx::execute_kernel([&](sclx::kernel_handler& handler) {
                handler.launch(
                    sclx::md_range_t<1>{out.shape()},
                    out,
                    [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                        out[idx] = src[index[idx]];
                    }
                );
            }).get();
        };
        step();
        return do_dump ? dump(out, path) : bench(step, steps, warmup);
    }
    if (name == "gather_x_selfcheck") {
        // config 4 at any size without a host copy of the result: the framework-owned gather
        // (binned form) against the user-lambda gather on the device; output = {mismatches, n}
#ifdef DRIVER_HAS_GATHER
        sclx::array<float, 1> src({4 * n}), out({n}), want({n});
        sclx::array<size_t, 1> index({n});
        const size_t m = 4 * n;
        generate(src, [] __device__(size_t i) { return uniform01(0xABC, i) + float(i >> 20); });
        generate(index, [=] __device__(size_t i) { return size_t(splitmix64(0x6A7, i) % m); });
        sclx::algorithm::gather(src, index, out, sclx::algorithm::gather_mode::exchange);
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{want.shape()},
                want,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) { want[idx] = src[index[idx]]; }
            );
        }).get();
        sclx::array<unsigned long long, 1> differs({n});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{differs.shape()},
                differs,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    differs[idx] = __float_as_uint(out[idx]) != __float_as_uint(want[idx]);
                }
            );
        }).get();
        const unsigned long long bad = sclx::algorithm::reduce(differs, 0ull, sclx::algorithm::plus<>());
        std::vector<unsigned long long> res{bad, (unsigned long long)n};
        if (!do_dump) return 2;
        return write_file(path, res.data(), res.size() * sizeof(unsigned long long));
#else
        return 2;
#endif
    }
    if (name == "gather_x_skew") return gather_x<float>(do_dump, n, path, steps, warmup, true);
    if (name == "gather_x_f32" || name == "gather_x_f64") {
        // configs[3] through the framework-owned gather (exchange form across partitions); the
        // reference build runs the same gather as the user-lambda idiom it documents
        return name == "gather_x_f32" ? gather_x<float>(do_dump, n, path, steps, warmup)
                                      : gather_x<double>(do_dump, n, path, steps, warmup);
    }
    if (name == "fill_iota") {
        // fill a rank-3 array, iota a rank-1 array, add them element-wise through a lambda
        sclx::array<int, 3> f({4, 4, n / 16});
        sclx::array<int, 1> io({n});
        auto step = [&] {
            sclx::fill(f, 41);
            sclx::iota(io, 1);
        };
        step();
        sclx::array<int, 1> r({n});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{r.shape()},
                r,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    r[idx] = f[idx[0]] + io[idx];
                }
            );
        }).get();
        return do_dump ? dump(r, path) : bench(step, steps, warmup);
    }
#ifdef DRIVER_HAS_CEREAL
    if (name == "serialize") {
        // serialize_array / deserialize_array round trip through cereal's binary archives
        // (scalix/array.cuh:907-950), then a kernel over the restored array
        sclx::array<int, 3> a({3, 4, n / 12});
        generate(a, [] __device__(size_t i) { return int(splitmix64(0x5E7, i)); });
        std::stringstream ss;
        {
            cereal::BinaryOutputArchive oar(ss);
            sclx::serialize_array(oar, a);
        }
        sclx::array<int, 3> b;
        {
            cereal::BinaryInputArchive iar(ss);
            sclx::deserialize_array(iar, b);
        }
        if (b.shape() != a.shape()) return 5;
        sclx::array<int, 3> r(a.shape());
        auto step = [&] {
            sclx::execute_kernel([&](sclx::kernel_handler& handler) {
                handler.launch(
                    sclx::md_range_t<3>{r.shape()},
                    r,
                    [=] __device__(const sclx::md_index_t<3>& idx, const auto&) { r[idx] = b[idx] + 1; }
                );
            }).get();
        };
        step();
        return do_dump ? dump(r, path) : bench(step, steps, warmup);
    }
#endif
    if (name == "algos") {
        // the thin algorithm/ entry points: transform, elementwise_reduce,
        // assign_array (int -> float), count_if, min_element, max_element, reduce.
        // Output: n values t[i] + e[i] + int(f[i]), then count, min, max, sum(t).
        sclx::array<int, 1> a({n}), b({n}), c({n}), t({n}), e({n});
        sclx::array<float, 1> f({n});
        generate(a, [] __device__(size_t i) { return int(splitmix64(0xA1, i) & 0xFFFF); });
        generate(b, [] __device__(size_t i) { return int(splitmix64(0xB2, i) & 0xFF); });
        generate(c, [] __device__(size_t i) { return int(splitmix64(0xC3, i) & 0xF); });
        sclx::algorithm::transform(t, a, 3, sclx::algorithm::multiplies<>()).get();
        sclx::algorithm::elementwise_reduce(sclx::algorithm::plus<>(), e, a, b, c).get();
        sclx::assign_array(a, f).get();
        const size_t evens = sclx::algorithm::count_if(a, is_even{});
        const int lo       = sclx::algorithm::min_element(a);
        const int hi       = sclx::algorithm::max_element(a);
        const int total    = sclx::algorithm::reduce(t, 0, sclx::algorithm::plus<>());
        sclx::array<int, 1> r({n});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{r.shape()},
                r,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    r[idx] = t[idx] + e[idx] + int(f[idx]);
                }
            );
        }).get();
        if (!do_dump) return 2;
        std::vector<int> host(n + 4);
        cudaDeviceSynchronize();
        if (cudaMemcpy(host.data(), r.data().get(), sizeof(int) * n, cudaMemcpyDefault) != cudaSuccess) return 3;
        host[n]     = int(evens);
        host[n + 1] = lo;
        host[n + 2] = hi;
        host[n + 3] = total;
        FILE* out = std::fopen(path, "wb");
        if (!out) return 4;
        std::fwrite(host.data(), sizeof(int), host.size(), out);
        std::fclose(out);
        return 0;
    }
    if (name == "errors") {
        // error behaviour of the API: which exception class each misuse raises.
        // One int per probe: 0 = no throw, 1 = std::invalid_argument, 2 = std::runtime_error,
        // 3 = any other std::exception.  `n` is the size of the arrays involved.
        auto probe = [](auto&& body) -> int {
            try {
                body();
            } catch (const std::invalid_argument&) {
                return 1;
            } catch (const std::runtime_error&) {
                return 2;
            } catch (const std::exception&) {
                return 3;
            }
            return 0;
        };
        std::vector<int> codes;
        // shape initialiser list of the wrong length (scalix/shape.cuh:66-69)
        codes.push_back(probe([&] { sclx::shape_t<2> s{1, 2, 3}; (void)s; }));
        // inclusive_scan into a result of another shape (scalix/algorithm/inclusive_scan.cuh:52-55)
        codes.push_back(probe([&] {
            sclx::array<int, 1> a({n}), r({n + 1});
            sclx::algorithm::inclusive_scan(a, r, 0, sclx::algorithm::plus<>());
        }));
        // reduce_last_dim into a result of the wrong shape (scalix/algorithm/reduce_last_dim.cuh:171-174)
        codes.push_back(probe([&] {
            sclx::array<uint, 2> a({3, n});
            sclx::array<uint, 1> r({4});
            sclx::algorithm::reduce_last_dim(a, r, 0u, sclx::algorithm::plus<>());
        }));
        // assign_array between arrays of different sizes (scalix/assign_array.cuh:42-45)
        codes.push_back(probe([&] {
            sclx::array<int, 1> a({n}), b({n / 2});
            sclx::assign_array(a, b).get();
        }));
        // re-partitioning a slice (scalix/array.cuh:525-529)
        codes.push_back(probe([&] {
            sclx::array<int, 2> a({4, n});
            auto row = a.get_slice(sclx::md_index_t<1>{{1}});
            row.set_primary_devices(std::vector<int>{0});
        }));
        // device id beyond the device count (scalix/array.cuh:491-499)
        codes.push_back(probe([&] {
            sclx::array<int, 1> a({n});
            a.set_primary_devices(std::vector<int>{0, 4096});
        }));
        // an exception raised inside the host lambda reaches the caller through future::get()
        // (scalix/execute_kernel.cuh:506-513)
        codes.push_back(probe([&] {
            sclx::execute_kernel([&](sclx::kernel_handler&) { throw std::runtime_error("from the host lambda"); }).get();
        }));
        // control: well-formed calls do not throw, and the runtime still works afterwards
        codes.push_back(probe([&] {
            sclx::array<int, 1> a({n}), r({n});
            sclx::fill(a, 1);
            sclx::algorithm::inclusive_scan(a, r, 0, sclx::algorithm::plus<>());
            if (r[n - 1] != int(n)) throw std::logic_error("scan after the error probes is wrong");
        }));
        if (!do_dump) return 2;
        FILE* f = std::fopen(path, "wb");
        if (!f) return 3;
        std::fwrite(codes.data(), sizeof(int), codes.size(), f);
        std::fclose(f);
        return 0;
    }
    if (name == "concurrent") {
        // two host threads run multi-partition algorithms at the same time (legal in the
        // reference: its partials and mutexes are function-local,
        // scalix/algorithm/reduce_last_dim.cuh:180-211, inclusive_scan.cuh:75-123); every result
        // must equal the one the same thread's data gives when run alone.  Output: one int per
        // (thread, round, algorithm): 0 = equal, else the number of differing values.
        constexpr int kThreads = 2, kRounds = 6;
        struct work_t {
            sclx::array<unsigned, 2> m;
            sclx::array<long long, 1> v;
            std::vector<unsigned> rows;
            std::vector<long long> scanned;
            long long total = 0;
        };
        std::vector<work_t> work(kThreads);
        for (int t = 0; t < kThreads; ++t) {
            work[t].m = sclx::array<unsigned, 2>({48, n});
            work[t].v = sclx::array<long long, 1>({4 * n + 3});
            auto m = work[t].m;
            auto v = work[t].v;
            const unsigned long long seed = 0xC0 + t;
            generate(m, [=] __device__(size_t i) { return unsigned(splitmix64(seed, i) & 0xFFFFF); });
            generate(v, [=] __device__(size_t i) { return (long long)(splitmix64(seed + 16, i) & 0xFFFF) - 100; });
            auto r = sclx::algorithm::reduce_last_dim(m, 0u, sclx::algorithm::plus<>());
            auto s = sclx::algorithm::inclusive_scan(v, 0ll, sclx::algorithm::plus<>());
            work[t].total = sclx::algorithm::reduce(v, 0ll, sclx::algorithm::plus<>());
            work[t].rows.assign(r.begin(), r.end());
            work[t].scanned.assign(s.begin(), s.end());
        }
        std::vector<int> verdict(size_t(kThreads) * kRounds * 3, -1);
        std::atomic<int> ready{0};
        std::vector<std::thread> threads;
        for (int t = 0; t < kThreads; ++t)
            threads.emplace_back([&, t] {
                ready.fetch_add(1);
                while (ready.load() < kThreads) {}  // start together
                try {
                    for (int round = 0; round < kRounds; ++round) {
                        int* out = &verdict[(size_t(t) * kRounds + round) * 3];
                        auto r   = sclx::algorithm::reduce_last_dim(work[t].m, 0u, sclx::algorithm::plus<>());
                        int bad  = 0;
                        for (size_t k = 0; k < work[t].rows.size(); ++k) bad += r[k] != work[t].rows[k];
                        out[0]   = bad;
                        auto s   = sclx::algorithm::inclusive_scan(work[t].v, 0ll, sclx::algorithm::plus<>());
                        bad      = 0;
                        for (size_t k = 0; k < work[t].scanned.size(); ++k) bad += s[k] != work[t].scanned[k];
                        out[1]   = bad;
                        out[2]   = sclx::algorithm::reduce(work[t].v, 0ll, sclx::algorithm::plus<>()) != work[t].total;
                    }
                } catch (const std::exception& e) {
                    std::fprintf(stderr, "concurrent: thread %d: %s\n", t, e.what());
                }
            });
        for (auto& th : threads) th.join();
        if (!do_dump) return 2;
        return write_file(path, verdict.data(), verdict.size() * sizeof(int));
    }
    if (name == "reshape_flatten") {
        // reshape / flatten on the host are COPIES with a new shape (scalix/array.cuh:882-905):
        // changing the source afterwards must not show in them
        sclx::array<int, 3> a({3, 4, n / 12});
        generate(a, [] __device__(size_t i) { return int(splitmix64(0x4E5, i) & 0xFFFFFF); });
        auto r2 = sclx::reshape(a, sclx::shape_t<2>{12, n / 12});
        auto f1 = sclx::flatten(a);
        sclx::fill(a, -1);
        int bad_shape = (r2.shape()[0] != 12) + (r2.shape()[1] != n / 12) + (f1.elements() != a.elements());
        int threw     = 0;
        try {
            auto wrong = sclx::reshape(a, sclx::shape_t<2>{5, n});
            (void)wrong;
        } catch (const std::invalid_argument&) {
            threw = 1;
        }
        sclx::array<int, 1> out({a.elements() + 2});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{out.shape()},
                out,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    const size_t i = idx[0];
                    if (i < f1.elements()) out[idx] = f1[i] + r2(i % 12, i / 12) + a[i];  // 2*value - 1
                    else out[idx] = i == f1.elements() ? bad_shape : threw;
                }
            );
        }).get();
        if (!do_dump) return 2;
        return dump(out, path);
    }
    if (name == "prefetch") {
        // both prefetch_async overloads (scalix/array.cuh:647-711), host reads and writes after a
        // prefetch to the CPU, and a read(2) from a file straight into array memory
        sclx::array<int, 2> a({4, n});
        generate(a, [] __device__(size_t i) { return int(i & 0x7FFFFFF); });
        a.prefetch_async({sclx::cuda::traits::cpu_device_id}).get();
        long long host_sum = 0;
        for (size_t i = 0; i < a.elements(); ++i) host_sum += a[i];
        for (size_t i = 0; i < a.elements(); i += 7) a[i] += 1000;
        // a file whose contents land in the second half of the array through the kernel's read(2)
        const size_t half = a.elements() / 2;
        std::vector<int> disk(half);
        for (size_t i = 0; i < half; ++i) disk[i] = int(3 * i + 1);
        const std::string tmp = std::string(path ? path : "/tmp/prefetch_case") + ".tmp";
        if (write_file(tmp.c_str(), disk.data(), disk.size() * sizeof(int))) return 4;
        const int fd = ::open(tmp.c_str(), O_RDONLY);
        if (fd < 0) return 4;
        size_t got = 0;
        char* dst  = reinterpret_cast<char*>(a.data().get() + half);
        while (got < half * sizeof(int)) {
            const ssize_t k = ::read(fd, dst + got, half * sizeof(int) - got);
            if (k <= 0) {
                std::fprintf(stderr, "prefetch: read(2) into array memory failed (errno %d)\n", errno);
                ::close(fd);
                return 6;
            }
            got += size_t(k);
        }
        ::close(fd);
        ::unlink(tmp.c_str());
        a.prefetch_async(sclx::exec_topology::distributed).get();
        a.prefetch_async().get();
        sclx::array<long long, 1> out({a.elements() + 1});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{out.shape()},
                out,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    const size_t i = idx[0];
                    out[idx]       = i < a.elements() ? 2ll * a[i] : host_sum;
                }
            );
        }).get();
        if (!do_dump) return 2;
        return dump(out, path);
    }
    if (name == "copy_range") {
        // array::copy_range_from (scalix/array.cuh:406-420): host data into columns [c0, c1)
        sclx::array<float, 2> a({4, n});
        sclx::fill(a, -2.0f);
        const size_t c0 = n / 3, c1 = n - n / 5;
        std::vector<float> host(4 * (c1 - c0));
        for (size_t i = 0; i < host.size(); ++i) host[i] = float(i % 4099) * 0.5f;
        a.copy_range_from(sclx::md_index_t<2>{0, c0}, sclx::md_index_t<2>{0, c1}, host.data());
        // a second piece from DEVICE memory
        sclx::array<float, 1> piece({8});
        sclx::fill(piece, 7.25f);
        a.copy_range_from(
            sclx::md_index_t<2>{0, 0}, sclx::md_index_t<2>{0, 2}, piece.data().get(), sclx::copy_policy::devicedevice
        );
        if (!do_dump) return 2;
        return dump(a, path);
    }
    if (name == "tuple_launch") {
        // launch over an array_tuple of arrays with different value types and ranks, range and
        // index-generator forms (scalix/execute_kernel.cuh:195-239)
        sclx::array<int, 1> ti({n});
        sclx::array<float, 2> tf({2, n});
        sclx::array_tuple<sclx::array<int, 1>, sclx::array<float, 2>> results(ti, tf);
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{ti.shape()},
                results,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    ti[idx]        = int(idx[0] * 3);
                    tf(0, idx[0])  = float(idx[0]) * 0.25f;
                    tf(1, idx[0])  = -float(idx[0]);
                }
            );
        }).get();
        sclx::array<int, 1> hits({n});
        sclx::array<float, 1> weight({n});
        sclx::array_tuple<sclx::array<int, 1>, sclx::array<float, 1>> results2(hits, weight);
        hash_index_generator<1, 1> gen(sclx::shape_t<1>{3 * n}, hits.shape(), 0x7B1E);
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                gen,
                results2,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    atomicAdd(&hits[idx], 1);
                    atomicAdd(&weight[idx], 0.5f);
                }
            );
        }).get();
        bool threw = false;
        try {
            sclx::array<int, 1> shorter({n / 2});
            sclx::array_tuple<sclx::array<int, 1>, sclx::array<int, 1>> bad(ti, shorter);
            sclx::execute_kernel([&](sclx::kernel_handler& handler) {
                handler.launch(sclx::md_range_t<1>{ti.shape()}, bad, [=] __device__(const sclx::md_index_t<1>&, const auto&) {});
            }).get();
        } catch (const std::invalid_argument&) {
            threw = true;
        }
        sclx::array<float, 1> out({n + 1});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{out.shape()},
                out,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                    const size_t i = idx[0];
                    out[idx] = i < n ? float(ti[i]) + tf(0, i) + tf(1, i) + float(hits[i]) * 16.f + weight[i]
                                     : (threw ? 1.f : 0.f);
                }
            );
        }).get();
        if (!do_dump) return 2;
        return dump(out, path);
    }
    if (name == "task_scheduler") {
        // cuda::task_scheduler::submit_task (scalix/cuda.hpp:270-303): the task runs on a thread
        // bound to `device_id`; results and exceptions come back through the future
        const int ndev = sclx::cuda::traits::device_count();
        std::vector<int> out;
        for (int d = 0; d < ndev; ++d) {
            auto fut = sclx::cuda::task_scheduler::submit_task(
                d,
                [](int a, int b) {
                    int* scratch = nullptr;
                    int host     = 0;
                    if (cudaMalloc(&scratch, sizeof(int)) != cudaSuccess) return -1;  // needs a current device
                    cudaMemset(scratch, 0, sizeof(int));
                    cudaMemcpy(&host, scratch, sizeof(int), cudaMemcpyDeviceToHost);
                    cudaFree(scratch);
                    return a * b + host;
                },
                6, 7
            );
            out.push_back(fut.get());
        }
        int caught = 0;
        try {
            sclx::cuda::task_scheduler::submit_task(0, [] { throw std::runtime_error("task"); }).get();
        } catch (const std::runtime_error&) {
            caught = 1;
        }
        out.resize(n, 42);  // independent of the device count: n values
        out[n - 1] = caught;
        if (!do_dump) return 2;
        return write_file(path, out.data(), out.size() * sizeof(int));
    }
    if (name == "future") {
        // the future of execute_kernel (scalix/execute_kernel.cuh:506-513): never deferred,
        // pollable, and DROPPING it leaves nothing running — a Thrust call on the raw pointers
        // from the default stream right after must see the results
        // (examples/index_generator.cu:78-88, examples/distributed_indexing.cu:94-105)
        sclx::array<int, 1> a({n}), b({n});
        std::vector<int> out;
        auto launch_into = [&](sclx::array<int, 1>& dst, int value, int spin) {
            return sclx::execute_kernel([&](sclx::kernel_handler& handler) {
                handler.launch(
                    sclx::md_range_t<1>{dst.shape()},
                    dst,
                    [=] __device__(const sclx::md_index_t<1>& idx, const auto&) {
                        int v = value;
                        for (int k = 0; k < spin; ++k) v = (v * 1664525 + 1013904223) | 1;  // keeps the kernel busy
                        dst[idx] = spin ? (v & 0) + value : value;
                    }
                );
            });
        };
        {
            auto fut = launch_into(a, 5, 0);
            out.push_back(fut.wait_for(std::chrono::seconds(0)) == std::future_status::deferred);  // 0
            int polls = 0;
            while (fut.wait_for(std::chrono::milliseconds(1)) != std::future_status::ready && polls < 100000) ++polls;
            out.push_back(polls < 100000);  // 1: polling terminates
            fut.get();
        }
        {
            auto fut = launch_into(b, 9, 20000);  // long kernel
            out.push_back(fut.wait_for(std::chrono::seconds(0)) == std::future_status::deferred);  // 0
            out.push_back(fut.valid());
            fut.get();
        }
        launch_into(a, 3, 20000);  // future dropped: must be complete when the statement ends
        out.push_back(thrust::reduce(thrust::device, a.begin(), a.end(), 0) == int(3 * n));  // 1
        launch_into(b, 4, 0);
        out.push_back(thrust::reduce(thrust::device, b.begin(), b.end(), 0) == int(4 * n));  // 1
        if (!do_dump) return 2;
        return write_file(path, out.data(), out.size() * sizeof(int));
    }
    if (name == "gen_hist") {
        // index-generator launch (scalix/execute_kernel.cuh:122-193; the pattern of
        // examples/index_generator.cu:72-93): every partition walks the whole generator range and
        // increments the bins it owns; the histogram must match bin for bin
        sclx::array<int, 3> bins({4, 4, n});
        sclx::fill(bins, 0);
        hash_index_generator<2, 3> gen(sclx::shape_t<2>{100, 100}, bins.shape(), 0x91D);
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                gen,
                bins,
                [=] __device__(const sclx::md_index_t<3>& idx, const auto&) { atomicAdd(&bins[idx], 1); }
            );
        }).get();
        if (!do_dump) return 2;
        return dump(bins, path);
    }
    if (name == "scaling_perf") {
        // the kernel of examples/scaling_performance.cu:59-78 (random read through a generator,
        // a local_array cache, 1000 x sqrt): every target element must be sqrtf(2)
        sclx::array<float, 3> source({4, 4, n}), target({4, 4, n});
        sclx::fill(source, 2.0f);
        hash_index_generator<3, 3> generator(target.shape(), source.shape(), 0x5CA1);
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            sclx::local_array<float, 1> local_cache(handler, sclx::cuda::traits::kernel::default_block_shape);
            handler.launch(
                sclx::md_range_t<3>{target.shape()},
                target,
                [=] __device__(const sclx::md_index_t<3>& index, const auto& info) mutable {
                    auto local_thread_id         = info.local_thread_id();
                    local_cache[local_thread_id] = source[generator(index)];
                    for (int i = 0; i < 1000; ++i) target[index] = sqrt(local_cache[local_thread_id]);
                }
            );
        }).get();
        if (!do_dump) return 2;
        return dump(target, path);
    }
#ifdef DRIVER_HAS_CEREAL
    if (name == "serialize_file") {
        // serialize to a FILE and deserialize through an std::ifstream (scalix/array.cuh:907-950):
        // for a large payload libstdc++ hands the destination pointer straight to read(2)
        sclx::array<int, 2> a({16, n / 16});
        generate(a, [] __device__(size_t i) { return int(splitmix64(0xF11E, i)); });
        const std::string tmp = std::string(path ? path : "/tmp/serialize_file_case") + ".tmp";
        {
            std::ofstream os(tmp, std::ios::binary);
            cereal::BinaryOutputArchive oar(os);
            sclx::serialize_array(oar, a);
        }
        sclx::array<int, 2> b;
        {
            std::ifstream is(tmp, std::ios::binary);
            cereal::BinaryInputArchive iar(is);
            sclx::deserialize_array(iar, b);
        }
        ::unlink(tmp.c_str());
        if (b.shape() != a.shape()) return 5;
        // checksum on the device: sum of (a[i] == b[i]) and a mixed hash of b
        sclx::array<int, 1> same({a.elements()});
        sclx::execute_kernel([&](sclx::kernel_handler& handler) {
            handler.launch(
                sclx::md_range_t<1>{same.shape()},
                same,
                [=] __device__(const sclx::md_index_t<1>& idx, const auto&) { same[idx] = a[idx[0]] == b[idx[0]]; }
            );
        }).get();
        const int equal = sclx::algorithm::reduce(same, 0, sclx::algorithm::plus<>());
        std::vector<int> out{equal == int(a.elements()), int(a.elements() & 0x7FFFFFFF)};
        if (!do_dump) return 2;
        return write_file(path, out.data(), out.size() * sizeof(int));
    }
#endif
    std::fprintf(stderr, "unknown case %s\n", name.c_str());
    return 2;
}

int main(int argc, char** argv) {
    if (argc < 5) {
        std::fprintf(stderr, "usage: %s dump <case> <n> <outfile> | bench <case> <n> <steps> <warmup> | bench_e2e rld_f32 <n> <steps> <warmup>\n", argv[0]);
        return 2;
    }
    const std::string what = argv[1], name = argv[2];
    const size_t n = std::strtoull(argv[3], nullptr, 10);
    try {
        if (what == "dump") return run_case(what, name, n, argv[4], 0, 0);
        if (what == "bench") return run_case(what, name, n, nullptr, std::atoi(argv[4]), argc > 5 ? std::atoi(argv[5]) : 3);
        if (what == "bench_e2e" && name == "rld_f32") return bench_e2e_rld(n, std::atoi(argv[4]), argc > 5 ? std::atoi(argv[5]) : 1);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "parity_driver: %s\n", e.what());
        return 1;
    }
    return 2;
}
