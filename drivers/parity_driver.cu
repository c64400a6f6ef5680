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
// cases: scan_i32 scan_i64 rld_f32 rld_u32max axpy_f32 axpy_f32_r3 gather_f32 fill_iota serialize gather_x_f32 gather_x_f64 gather_x_skew errors algos
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

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

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

template<class Step>
int bench(Step&& step, int steps, int warmup) {
    for (int i = 0; i < warmup; ++i) step();
    cudaDeviceSynchronize();
    timer t;
    for (int i = 0; i < steps; ++i) step();
    cudaDeviceSynchronize();
    std::printf("{\"ms_per_step\": %.6f, \"steps\": %d, \"warmup\": %d}\n", t.ms() / steps, steps, warmup);
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
    std::fprintf(stderr, "unknown case %s\n", name.c_str());
    return 2;
}

int main(int argc, char** argv) {
    if (argc < 5) {
        std::fprintf(stderr, "usage: %s dump <case> <n> <outfile> | bench <case> <n> <steps> <warmup>\n", argv[0]);
        return 2;
    }
    const std::string what = argv[1], name = argv[2];
    const size_t n = std::strtoull(argv[3], nullptr, 10);
    try {
        if (what == "dump") return run_case(what, name, n, argv[4], 0, 0);
        if (what == "bench") return run_case(what, name, n, nullptr, std::atoi(argv[4]), argc > 5 ? std::atoi(argv[5]) : 3);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "parity_driver: %s\n", e.what());
        return 1;
    }
    return 2;
}
