// Experiment: how many DRAM bytes does one random 4-byte load cost on B200, and which knob
// changes it?  (ncu on the exchange gather's serve kernel showed ~123 B of DRAM reads per 4-byte
// gather: the random-read "rate limit" of 40 G/s is really 5.2 TB/s of over-fetched lines.)
// Variants: load flavour (plain nc, L2::64B/128B/256B prefetch hints, .cs/.cv/.lu, L2 evict_first
// policy, no-allocate) x cudaLimitMaxL2FetchGranularity (32/64/128).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e__)); exit(2); } } while (0)

__device__ __forceinline__ unsigned long long mix(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

enum { PLAIN, NC, H64, H128, H256, CS, CV, LU, EVICT_FIRST, NOALLOC, NC_H64 };

template<int KIND>
__device__ __forceinline__ unsigned load(const unsigned* p, unsigned long long pol) {
    unsigned v;
    if constexpr (KIND == PLAIN) asm volatile("ld.global.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == NC) asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == H64) asm volatile("ld.global.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == H128) asm volatile("ld.global.L2::128B.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == H256) asm volatile("ld.global.L2::256B.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == CS) asm volatile("ld.global.cs.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == CV) asm volatile("ld.global.cv.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == LU) asm volatile("ld.global.lu.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == EVICT_FIRST)
        asm volatile("ld.global.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    if constexpr (KIND == NOALLOC) asm volatile("ld.global.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(p));
    if constexpr (KIND == NC_H64) asm volatile("ld.global.nc.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

template<int KIND>
__global__ void rand_read(const unsigned* p, size_t n_elems, size_t n_req, unsigned* sink) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    unsigned long long pol = 0;
    if constexpr (KIND == EVICT_FIRST) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    unsigned acc = 0;
    for (; i < n_req; i += stride * 4) {
        unsigned v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = load<KIND>(p + (mix(i + u * stride) & (n_elems - 1)), pol);
#pragma unroll
        for (int u = 0; u < 4; ++u) acc += v[u];
    }
    if (acc == 0x12345678u) *sink = acc;
}

template<int KIND>
void run(const char* name, const unsigned* p, size_t n_elems, size_t n_req, unsigned* sink) {
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    rand_read<KIND><<<148 * 8, 256>>>(p, n_elems, n_req, sink); CK(cudaDeviceSynchronize());
    cudaEventRecord(s);
    for (int r = 0; r < 3; ++r) rand_read<KIND><<<148 * 8, 256>>>(p, n_elems, n_req, sink);
    cudaEventRecord(e); CK(cudaEventSynchronize(e));
    float ms; cudaEventElapsedTime(&ms, s, e); ms /= 3;
    printf("  %-12s %8.3f ms  %7.2f Gload/s\n", name, ms, n_req / ms / 1e6);
    fflush(stdout);
}

int main(int argc, char** argv) {
    const size_t bytes = size_t(4) << 30, n_elems = bytes / 4, n_req = size_t(1) << 27;
    unsigned *buf, *sink;
    CK(cudaMalloc(&buf, bytes)); CK(cudaMalloc(&sink, 4)); CK(cudaMemset(buf, 0, bytes));
    size_t lim = 0;
    CK(cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity));
    printf("default cudaLimitMaxL2FetchGranularity = %zu\n", lim);
    const size_t grans[] = {0, 32, 64, 128};
    for (size_t g : grans) {
        if (g) {
            cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g);
            CK(cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity));
            printf("set limit %zu -> %s, now %zu\n", g, cudaGetErrorString(e), lim);
        }
        run<PLAIN>("plain", buf, n_elems, n_req, sink);
        run<NC>("nc", buf, n_elems, n_req, sink);
        run<H64>("L2::64B", buf, n_elems, n_req, sink);
        run<H128>("L2::128B", buf, n_elems, n_req, sink);
        run<H256>("L2::256B", buf, n_elems, n_req, sink);
        run<NC_H64>("nc.L2::64B", buf, n_elems, n_req, sink);
        run<CS>("cs", buf, n_elems, n_req, sink);
        run<CV>("cv", buf, n_elems, n_req, sink);
        run<LU>("lu", buf, n_elems, n_req, sink);
        run<EVICT_FIRST>("evict_first", buf, n_elems, n_req, sink);
        run<NOALLOC>("no_allocate", buf, n_elems, n_req, sink);
    }
    return 0;
}
