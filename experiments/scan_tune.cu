// Experiment: sweep of scan tile geometries + CUB DeviceScan bars (reference path and tuned path).
#include "scalix/detail/b200/scan.cuh"
#include "scalix/detail/b200/reduce.cuh"
#include <cub/device/device_scan.cuh>
#include <thrust/functional.h>
#include <cuda/std/functional>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
using namespace sclx::b200;
#define CK(x) do { cudaError_t err__ = (x); if (err__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(err__)); exit(2); } } while (0)
struct plus_op { template<class T> __host__ __device__ T operator()(const T& a, const T& b) const { return a + b; } };

template<class T, int THREADS, int VECS>
float run_ours(const T* in, T* out, size_t n, void* desc, unsigned& epoch, int reps, const char* sel, const char* name) {
    if (sel && strcmp(sel, name)) return 0;
    using cfg = scan_config<T, THREADS, VECS>;
    size_t tiles = (n + cfg::tile_items - 1) / cfg::tile_items;
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    auto launch = [&] {
        inclusive_scan_kernel<T, plus_op, THREADS, VECS, true><<<unsigned(tiles), THREADS>>>(
            in, out, n, (typename tile_desc<T>::word_t*)desc, ++epoch, scan_carry_t{nullptr, 0}, nullptr, plus_op{});
    };
    for (int i = 0; i < 2; ++i) launch();
    CK(cudaDeviceSynchronize());
    cudaEventRecord(s);
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e); CK((cudaEventSynchronize(e)));
    float ms; cudaEventElapsedTime(&ms, s, e); ms /= reps;
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, inclusive_scan_kernel<T, plus_op, THREADS, VECS, true>, THREADS, 0);
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, inclusive_scan_kernel<T, plus_op, THREADS, VECS, true>);
    printf("%-22s T=%zuB  %8.4f ms  %8.1f GB/s  (regs %d, occ %d CTA/SM, tile %zu B)\n", name, sizeof(T), ms,
           2.0 * n * sizeof(T) / ms / 1e6, fa.numRegs, occ, cfg::tile_items * sizeof(T));
    return ms;
}

template<class T, class Op>
void run_cub(const T* in, T* out, size_t n, int reps, const char* name, Op op) {
    void* tmp = nullptr; size_t tb = 0;
    cub::DeviceScan::InclusiveScan(tmp, tb, in, out, op, n);
    CK(cudaMalloc(&tmp, tb));
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    for (int i = 0; i < 2; ++i) cub::DeviceScan::InclusiveScan(tmp, tb, in, out, op, n);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(s);
    for (int i = 0; i < reps; ++i) cub::DeviceScan::InclusiveScan(tmp, tb, in, out, op, n);
    cudaEventRecord(e); CK((cudaEventSynchronize(e)));
    float ms; cudaEventElapsedTime(&ms, s, e); ms /= reps;
    printf("%-22s T=%zuB  %8.4f ms  %8.1f GB/s\n", name, sizeof(T), ms, 2.0 * n * sizeof(T) / ms / 1e6);
    cudaFree(tmp);
}

template<class T>
void sweep(size_t n, const char* sel, int reps) {
    T *in, *out; void* desc;
    CK(cudaMalloc(&in, n * sizeof(T))); CK(cudaMalloc(&out, n * sizeof(T)));
    CK(cudaMalloc(&desc, (n / 512 + 64) * 16)); CK(cudaMemset(desc, 0, (n / 512 + 64) * 16));
    CK(cudaMemset(in, 1, n * sizeof(T)));
    unsigned epoch = 0;
#define R(TH, VE) run_ours<T, TH, VE>(in, out, n, desc, epoch, reps, sel, #TH "x" #VE)
    R(128, 4); R(128, 8); R(256, 2); R(256, 4); R(256, 6); R(256, 8); R(384, 4); R(384, 6); R(512, 2); R(512, 4); R(512, 6); R(1024, 2); R(1024, 4);
    if (!sel) {
        run_cub<T>(in, out, n, reps, "cub thrust::plus<>", thrust::plus<void>{});
        run_cub<T>(in, out, n, reps, "cub cuda::std::plus<T>", cuda::std::plus<T>{});
        // plain copy for the roofline denominator on this box
        cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
        cudaMemcpy(out, in, n * sizeof(T), cudaMemcpyDeviceToDevice);
        cudaEventRecord(s); for (int i = 0; i < reps; ++i) cudaMemcpyAsync(out, in, n * sizeof(T), cudaMemcpyDeviceToDevice);
        cudaEventRecord(e); cudaEventSynchronize(e); float ms; cudaEventElapsedTime(&ms, s, e); ms /= reps;
        printf("%-22s T=%zuB  %8.4f ms  %8.1f GB/s\n", "cudaMemcpy D2D", sizeof(T), ms, 2.0 * n * sizeof(T) / ms / 1e6);
    }
    cudaFree(in); cudaFree(out); cudaFree(desc);
}

int main(int argc, char** argv) {
    const char* sel = argc > 2 ? argv[2] : nullptr;     // e.g. "256x6"
    const char* ty  = argc > 1 ? argv[1] : "all";
    int reps = argc > 3 ? atoi(argv[3]) : 10;
    if (!strcmp(ty, "all") || !strcmp(ty, "i32")) sweep<unsigned>(size_t(1) << 29, sel, reps);
    if (!strcmp(ty, "all") || !strcmp(ty, "i64")) sweep<unsigned long long>(size_t(1) << 28, sel, reps);
    return 0;
}
