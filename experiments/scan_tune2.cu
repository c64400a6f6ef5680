// Experiment: geometry sweep of the streaming scan (scan_stream.cuh) against the register
// look-back kernel (scan.cuh), CUB DeviceScan (both functor flavours) and a D2D copy.
// Every configuration is checked element-for-element against the CUB result.
#ifdef STATS
#define SCLX_SCAN_STATS 1
#endif
#include "scalix/detail/b200/scan_stream.cuh"
#include <cub/device/device_scan.cuh>
#include <thrust/functional.h>
#include <cuda/std/functional>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>
using namespace sclx::b200;
#define CK(x) do { cudaError_t err__ = (x); if (err__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(err__)); exit(2); } } while (0)
struct plus_op { template<class T> __host__ __device__ T operator()(const T& a, const T& b) const { return a + b; } };

template<class T>
__global__ void init_kernel(T* p, size_t n) {
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        unsigned long long z = i * 0x9E3779B97F4A7C15ull; z ^= z >> 29;
        p[i] = T(z & 0xFFFF);
    }
}
template<class T>
__global__ void diff_kernel(const T* a, const T* b, size_t n, unsigned long long* bad) {
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        if (a[i] != b[i]) atomicAdd(bad, 1ull);
}

static int g_sms = 148;
static void* g_flush = nullptr;  // FLUSH=1: L2 eviction buffer

template<class T, int VECS, int CHUNKS, int STAGES, int WR, int WS, int LBK, int WLB, int RUN = 1>
void run_stream(const T* in, T* out, const T* want, size_t n, void* desc, unsigned& epoch, scan_stream_counters* ctr,
                unsigned long long* bad, int reps, const char* sel, int force_cps) {
    char name[64];
    if (RUN == 1) snprintf(name, sizeof name, "v%dc%ds%d/r%dw%dk%dl%d", VECS, CHUNKS, STAGES, WR, WS, LBK, WLB);
    else snprintf(name, sizeof name, "v%dc%ds%d/r%dw%dk%dl%dR%d", VECS, CHUNKS, STAGES, WR, WS, LBK, WLB, RUN);
    if (sel && strcmp(sel, name)) return;
    using cfg = scan_stream_config<T, VECS, CHUNKS, STAGES, WR, WS, WLB>;
    constexpr int THREADS = cfg::threads;
    auto kern = inclusive_scan_stream_kernel<T, plus_op, VECS, CHUNKS, STAGES, WR, WS, LBK, WLB, RUN>;
    if (cfg::smem_bytes > 227 * 1024) return;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cfg::smem_bytes)));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, cfg::smem_bytes));
    if (occ < 1) { printf("%-12s does not fit\n", name); return; }
    int cps = force_cps > 0 && force_cps < occ ? force_cps : occ;
    size_t tiles = (n + cfg::tile_items - 1) / cfg::tile_items;
    unsigned grid = unsigned(std::min<size_t>(tiles, size_t(g_sms) * cps));
    auto launch = [&] {
        kern<<<grid, THREADS, cfg::smem_bytes>>>(in, out, n, (typename tile_desc<T>::word_t*)desc, ++epoch,
                                                 scan_carry_t{nullptr, 0}, nullptr, ctr, plus_op{});
    };
    CK(cudaMemset(out, 0, n * sizeof(T)));
    launch();
    CK(cudaDeviceSynchronize());
    CK(cudaMemset(bad, 0, 8));
    diff_kernel<<<1184, 256>>>(out, want, n, bad);
    unsigned long long hb = 0;
    CK(cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost));
    launch();
    CK(cudaDeviceSynchronize());
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    cudaEventRecord(s);
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e); CK((cudaEventSynchronize(e)));
    float ms; cudaEventElapsedTime(&ms, s, e); ms /= reps;
    if (g_flush) {
        // config-1 style timing: evict in/out from the 126 MB L2 before every launch, time each
        // launch on its own, report the median
        std::vector<float> t;
        for (int i = 0; i < reps; ++i) {
            CK(cudaMemsetAsync(g_flush, i, size_t(256) << 20));
            cudaEventRecord(s); launch(); cudaEventRecord(e); CK(cudaEventSynchronize(e));
            float one; cudaEventElapsedTime(&one, s, e); t.push_back(one);
        }
        std::sort(t.begin(), t.end());
        ms = t[t.size() / 2];
    }
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, kern);
#ifdef STATS
    {
        unsigned long long z[16] = {0}, h[16];
        cudaMemcpyToSymbol(g_scan_stats, z, sizeof z);
        launch();
        CK(cudaDeviceSynchronize());
        cudaMemcpyFromSymbol(h, g_scan_stats, sizeof h);
        double t = double(h[7] ? h[7] : 1), per_cta = double(tiles) / grid;
        printf("    per tile: lookback %.0f cyc, %.2f rounds, %.1f re-polls | per CTA-tile waits: scanner %.0f reducer %.0f loader %.0f storer %.0f cyc\n",
               h[0] / t, h[1] / t, h[2] / t, h[3] / (per_cta * grid), h[4] / (per_cta * grid), h[5] / (per_cta * grid), h[6] / (per_cta * grid));
        printf("    stage timeline since load issue (cyc): full %.0f  aggregate %.0f  lb-start %.0f  prefixed %.0f  written %.0f  empty %.0f | period %.0f\n",
               h[8] / t, h[9] / t, h[10] / t, h[11] / t, h[12] / t, h[13] / t, ms * 1e-3 * 1.92e9 / per_cta);
    }
#endif
    printf("stream %-12s T=%zuB  %8.4f ms  %8.1f GB/s  (regs %d, %d CTA/SM of max %d, tile %zu B, smem %zu B) %s\n", name,
           sizeof(T), ms, 2.0 * n * sizeof(T) / ms / 1e6, fa.numRegs, cps, occ, cfg::tile_bytes, cfg::smem_bytes,
           hb ? "MISMATCH" : "ok");
    if (hb) printf("   !!! %llu mismatching elements\n", hb);
    fflush(stdout);
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
    printf("%-26s T=%zuB  %8.4f ms  %8.1f GB/s\n", name, sizeof(T), ms, 2.0 * n * sizeof(T) / ms / 1e6);
    cudaFree(tmp);
}

template<class T>
void sweep(size_t n, const char* sel, int reps, int cps) {
    T *in, *out, *want; void* desc; scan_stream_counters* ctr; unsigned long long* bad;
    CK(cudaMalloc(&in, n * sizeof(T))); CK(cudaMalloc(&out, n * sizeof(T))); CK(cudaMalloc(&want, n * sizeof(T)));
    CK(cudaMalloc(&desc, (n / 512 + 64) * 16)); CK(cudaMemset(desc, 0, (n / 512 + 64) * 16));
    CK(cudaMalloc(&ctr, 256)); CK(cudaMemset(ctr, 0, 256));
    CK(cudaMalloc(&bad, 8));
    init_kernel<<<1184, 256>>>(in, n);
    unsigned epoch = 0;
    {
        void* tmp = nullptr; size_t tb = 0;
        cub::DeviceScan::InclusiveScan(tmp, tb, in, want, cuda::std::plus<T>{}, n);
        CK(cudaMalloc(&tmp, tb));
        cub::DeviceScan::InclusiveScan(tmp, tb, in, want, cuda::std::plus<T>{}, n);
        CK(cudaDeviceSynchronize());
        cudaFree(tmp);
    }
#define R(VE, CH, ST, WR, WS, K, L) run_stream<T, VE, CH, ST, WR, WS, K, L>(in, out, want, n, desc, epoch, ctr, bad, reps, sel, cps)
#ifdef SWEEP2
    // second sweep: LB_K = 1 on the geometries that were only tried with LB_K = 2, fewer / more
    // reducer and scanner warps around the two production points
    R(5, 16, 5, 8, 16, 1, 1); R(9, 8, 6, 8, 16, 1, 1); R(9, 8, 6, 8, 16, 2, 1); R(5, 14, 6, 8, 16, 1, 1); R(3, 24, 6, 8, 16, 1, 1);
    R(7, 12, 5, 8, 16, 1, 1); R(7, 10, 6, 8, 16, 1, 1); R(9, 8, 6, 4, 16, 1, 1); R(9, 8, 6, 8, 8, 1, 1); R(5, 16, 5, 4, 16, 1, 1);
    R(5, 16, 5, 8, 8, 1, 1); R(5, 16, 5, 16, 8, 1, 1); R(11, 6, 6, 6, 12, 1, 1); R(13, 5, 6, 5, 10, 1, 1); R(5, 12, 7, 8, 12, 1, 1);
    return;
#endif
#ifdef SWEEP4
    // config 1 (2^24 int32, FLUSH=1): does a smaller tile / deeper ring fill and drain faster?
    R(9, 8, 6, 4, 16, 1, 1); R(5, 16, 5, 8, 16, 1, 1); R(9, 4, 12, 4, 16, 1, 1); R(5, 8, 10, 4, 16, 1, 1); R(9, 6, 8, 4, 16, 1, 1);
    R(5, 4, 20, 4, 8, 1, 1); R(9, 2, 24, 2, 8, 1, 1); R(9, 8, 6, 8, 16, 1, 1); R(9, 8, 6, 4, 8, 1, 1); R(3, 8, 16, 4, 16, 1, 1);
    R(9, 3, 16, 3, 12, 1, 1); R(5, 6, 14, 6, 12, 1, 1);
    return;
#endif
#define RR(VE, CH, ST, WR, WS, K, L, RN) run_stream<T, VE, CH, ST, WR, WS, K, L, RN>(in, out, want, n, desc, epoch, ctr, bad, reps, sel, cps)
#ifdef SWEEP7
    // runs, taken only when RUN stages are free
    R(9, 8, 6, 4, 16, 1, 1); RR(9, 8, 6, 4, 16, 1, 1, 2); RR(9, 6, 8, 4, 16, 1, 1, 2); RR(9, 4, 12, 4, 16, 1, 1, 2);
    RR(9, 4, 12, 4, 16, 1, 1, 3); RR(9, 4, 12, 4, 16, 1, 1, 4); RR(9, 3, 16, 3, 12, 1, 1, 4); RR(5, 8, 10, 4, 16, 1, 1, 2);
    RR(9, 6, 8, 4, 16, 1, 1, 4); RR(9, 2, 24, 2, 8, 1, 1, 4); RR(9, 2, 24, 2, 8, 1, 1, 8); RR(9, 4, 12, 4, 8, 1, 1, 2);
    return;
#endif
#ifdef SWEEP6
    // runs of consecutive tiles per ticket: one look-back per run
    R(9, 8, 6, 4, 16, 1, 1); RR(9, 8, 6, 4, 16, 1, 1, 2); RR(9, 8, 6, 4, 16, 1, 1, 3); RR(9, 8, 6, 4, 16, 1, 1, 6);
    RR(9, 6, 8, 4, 16, 1, 1, 2); RR(9, 6, 8, 4, 16, 1, 1, 4); RR(9, 4, 12, 4, 16, 1, 1, 4); RR(5, 8, 10, 4, 16, 1, 1, 4);
    RR(9, 8, 6, 4, 16, 2, 1, 2); RR(9, 8, 6, 8, 16, 1, 1, 2); RR(9, 8, 6, 4, 12, 1, 1, 2); RR(9, 7, 7, 4, 16, 1, 1, 2);
    return;
#endif
#ifdef SWEEP5
    // look-back window width vs problem size: at start-up no tile carries a PREFIX yet, so the
    // prefix wavefront advances 32*LB_K tiles per L2 round trip
    R(9, 8, 6, 4, 16, 1, 1); R(9, 8, 6, 4, 16, 2, 1); R(9, 8, 6, 4, 16, 4, 1); R(9, 8, 6, 4, 16, 8, 1);
    return;
#endif
#ifdef SWEEP3
    // third sweep: around the sweep-2 winner v9c8s6/r4w16k1
    R(9, 8, 6, 4, 16, 1, 1); R(9, 8, 6, 2, 16, 1, 1); R(9, 8, 6, 3, 16, 1, 1); R(9, 8, 6, 4, 12, 1, 1); R(9, 8, 6, 4, 20, 1, 1);
    R(9, 8, 6, 4, 8, 1, 1); R(9, 8, 6, 2, 8, 1, 1); R(9, 8, 6, 4, 16, 1, 2); R(7, 10, 6, 4, 16, 1, 1); R(3, 24, 6, 4, 16, 1, 1);
    R(9, 6, 8, 4, 16, 1, 1); R(9, 10, 4, 4, 16, 1, 1); R(9, 12, 4, 4, 16, 1, 1); R(11, 7, 5, 4, 16, 1, 1); R(9, 9, 5, 4, 16, 1, 1);
    R(9, 7, 7, 4, 16, 1, 1); R(5, 14, 6, 4, 16, 1, 1); R(9, 8, 6, 4, 24, 1, 1);
    return;
#endif
    R(5, 16, 5, 8, 16, 1, 1); R(5, 16, 5, 8, 16, 2, 1); R(5, 16, 5, 8, 16, 3, 1); R(5, 16, 5, 8, 20, 2, 1); R(5, 16, 5, 6, 12, 2, 1); R(5, 16, 5, 12, 16, 2, 1);
    R(9, 8, 6, 8, 16, 2, 1); R(5, 8, 10, 8, 16, 2, 1); R(5, 12, 7, 8, 12, 2, 1); R(3, 32, 4, 8, 16, 2, 1); R(7, 16, 3, 8, 16, 2, 1); R(7, 12, 5, 8, 16, 2, 1);
    R(5, 20, 4, 8, 16, 2, 1); R(5, 18, 4, 8, 16, 2, 1); R(5, 14, 6, 8, 16, 2, 1); R(3, 24, 6, 8, 16, 2, 1); R(5, 22, 4, 8, 16, 2, 1);
    if (!sel) {
        run_cub<T>(in, out, n, reps, "cub thrust::plus<>", thrust::plus<void>{});
        run_cub<T>(in, out, n, reps, "cub cuda::std::plus<T>", cuda::std::plus<T>{});
        cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
        cudaMemcpy(out, in, n * sizeof(T), cudaMemcpyDeviceToDevice);
        cudaEventRecord(s); for (int i = 0; i < reps; ++i) cudaMemcpyAsync(out, in, n * sizeof(T), cudaMemcpyDeviceToDevice);
        cudaEventRecord(e); cudaEventSynchronize(e); float ms; cudaEventElapsedTime(&ms, s, e); ms /= reps;
        printf("%-26s T=%zuB  %8.4f ms  %8.1f GB/s\n", "cudaMemcpy D2D", sizeof(T), ms, 2.0 * n * sizeof(T) / ms / 1e6);
    }
    cudaFree(in); cudaFree(out); cudaFree(want); cudaFree(desc); cudaFree(ctr); cudaFree(bad);
}

int main(int argc, char** argv) {
    const char* ty  = argc > 1 ? argv[1] : "all";
    const char* sel = argc > 2 && strcmp(argv[2], "-") ? argv[2] : nullptr;  // e.g. "512x5x3"
    int reps = argc > 3 ? atoi(argv[3]) : 10;
    int cps  = argc > 4 ? atoi(argv[4]) : 0;                                  // cap CTAs/SM
    size_t lg = argc > 5 ? atoi(argv[5]) : 0;
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, 0);
    if (getenv("FLUSH")) CK(cudaMalloc(&g_flush, size_t(256) << 20));
    if (!strcmp(ty, "all") || !strcmp(ty, "i32")) sweep<unsigned>(size_t(1) << (lg ? lg : 29), sel, reps, cps);
    if (!strcmp(ty, "all") || !strcmp(ty, "i64")) sweep<unsigned long long>(size_t(1) << (lg ? lg : 28), sel, reps, cps);
    return 0;
}
