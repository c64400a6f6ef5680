// Experiment: geometry of the exchange gather's serve kernel (K2): resp[i] = src[queue[i]],
// queue = 2^27 random 32-bit offsets into a 2 GiB partition.  Sweeps CTAs/SM, unroll and the
// L2 prefetch-size hint of the random load.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e__)); exit(2); } } while (0)
__device__ __forceinline__ unsigned long long mix(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31);
}
__global__ void fill_queue(unsigned* q, size_t n, unsigned mask) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, s = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += s) q[i] = unsigned(mix(i)) & mask;
}
template<int HINT> __device__ __forceinline__ unsigned ld(const unsigned* p) {
    unsigned v;
    if constexpr (HINT == 0) asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (HINT == 64) asm volatile("ld.global.nc.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (HINT == 1) asm volatile("ld.global.nc.L1::no_allocate.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
template<int THREADS, int UNROLL, int HINT>
__global__ void __launch_bounds__(THREADS) serve(const unsigned* __restrict__ queue, size_t cnt, const unsigned* __restrict__ src, unsigned* __restrict__ resp) {
    const size_t gtid = size_t(blockIdx.x) * THREADS + threadIdx.x, gstride = size_t(gridDim.x) * THREADS;
    size_t i = gtid;
    for (; i + (UNROLL - 1) * gstride < cnt; i += UNROLL * gstride) {
        unsigned o[UNROLL], v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) o[u] = __ldcs(queue + i + u * gstride);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ld<HINT>(src + o[u]);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) resp[i + u * gstride] = v[u];
    }
    for (; i < cnt; i += gstride) resp[i] = ld<HINT>(src + __ldcs(queue + i));
}
template<int THREADS, int UNROLL, int HINT>
void run(const unsigned* q, size_t n, const unsigned* src, unsigned* resp, int per_sm) {
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    const int grid = 148 * per_sm;
    serve<THREADS, UNROLL, HINT><<<grid, THREADS>>>(q, n, src, resp); CK(cudaDeviceSynchronize());
    cudaEventRecord(s);
    for (int r = 0; r < 3; ++r) serve<THREADS, UNROLL, HINT><<<grid, THREADS>>>(q, n, src, resp);
    cudaEventRecord(e); CK(cudaEventSynchronize(e));
    float ms; cudaEventElapsedTime(&ms, s, e); ms /= 3;
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, serve<THREADS, UNROLL, HINT>, THREADS, 0);
    printf("threads %4d unroll %d hint %2d  %2d CTA/SM (fit %d)  %7.3f ms  %6.2f Ggather/s\n", THREADS, UNROLL, HINT, per_sm, occ, ms, n / ms / 1e6);
    fflush(stdout);
}
int main() {
    const size_t n = size_t(1) << 27, src_elems = size_t(1) << 29;
    unsigned *q, *src, *resp;
    CK(cudaMalloc(&q, n * 4)); CK(cudaMalloc(&src, src_elems * 4)); CK(cudaMalloc(&resp, n * 4));
    CK(cudaMemset(src, 1, src_elems * 4));
    fill_queue<<<148 * 8, 256>>>(q, n, unsigned(src_elems - 1)); CK(cudaDeviceSynchronize());
    if (getenv("SERVE_U1")) {  // one element per thread per trip: what an opaque lambda over the range launcher does
        for (int per_sm : {8, 16, 32, 64, 128, 512, 2048, 3543}) run<256, 1, 0>(q, n, src, resp, per_sm);
        return 0;
    }
    if (getenv("SERVE_WIDE")) {
        for (int per_sm : {16, 24, 32, 48, 64, 128, 256, 886}) run<256, 4, 64>(q, n, src, resp, per_sm);
        for (int per_sm : {16, 32, 64, 128}) run<256, 4, 0>(q, n, src, resp, per_sm);
        for (int per_sm : {32, 64, 128, 256}) run<256, 2, 64>(q, n, src, resp, per_sm);
        for (int per_sm : {32, 64, 128, 512}) run<128, 4, 64>(q, n, src, resp, per_sm);
        for (int per_sm : {16, 32, 64}) run<512, 4, 64>(q, n, src, resp, per_sm);
        return 0;
    }
    for (int per_sm : {2, 4, 6, 8, 12, 16}) run<256, 4, 0>(q, n, src, resp, per_sm);
    for (int per_sm : {4, 8, 16}) run<256, 1, 0>(q, n, src, resp, per_sm);
    for (int per_sm : {4, 8, 16}) run<256, 2, 0>(q, n, src, resp, per_sm);
    for (int per_sm : {2, 4, 8}) run<256, 8, 0>(q, n, src, resp, per_sm);
    for (int per_sm : {1, 2, 4}) run<1024, 4, 0>(q, n, src, resp, per_sm);
    for (int per_sm : {4, 6, 8, 16}) run<256, 4, 64>(q, n, src, resp, per_sm);
    for (int per_sm : {4, 8, 16}) run<256, 4, 1>(q, n, src, resp, per_sm);
    for (int per_sm : {4, 8, 16}) run<128, 4, 64>(q, n, src, resp, per_sm);
    return 0;
}
