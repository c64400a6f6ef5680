// Experiment: do the framework-owned streaming kernels (fill, iota, generate) also gain from
// more than one resident wave, like the range dispatcher does?
#include "scalix/detail/b200/stream_kernels.cuh"
#include <cstdio>
#include <cstdlib>
using namespace sclx::b200;
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e__)); exit(2); } } while (0)
struct copy_gen { const float* a; __device__ float operator()(size_t i) const { return a[i] * 2.0f; } };
template<class L> float timeit(L launch) {
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    launch(); launch(); CK(cudaDeviceSynchronize());
    cudaEventRecord(s); for (int r = 0; r < 5; ++r) launch(); cudaEventRecord(e); CK(cudaEventSynchronize(e));
    float ms; cudaEventElapsedTime(&ms, s, e); return ms / 5;
}
int main() {
    const size_t n = size_t(1) << 30;
    float *a, *b; CK(cudaMalloc(&a, n * 4)); CK(cudaMalloc(&b, n * 4)); CK(cudaMemset(a, 0, n * 4));
    for (int mult : {1, 2, 4, 8, 16, 64, 256}) {
        const unsigned grid = 148u * 8u * mult;
        float f = timeit([&] { fill_kernel<unsigned, 256, 4><<<grid, 256>>>((unsigned*)b, n, 7u); });
        float i = timeit([&] { iota_kernel<unsigned, 256, 4><<<grid, 256>>>((unsigned*)b, n, 1u, 0); });
        float g = timeit([&] { generate_kernel<float, copy_gen, 256, 8><<<grid, 256>>>(b, n, 0, copy_gen{a}); });
        printf("grid %3dx resident: fill %7.3f ms %7.1f GB/s | iota %7.3f ms %7.1f GB/s | generate(copy) %7.3f ms %7.1f GB/s\n",
               mult, f, 4.0 * n / f / 1e6, i, 4.0 * n / i / 1e6, g, 8.0 * n / g / 1e6);
        fflush(stdout);
    }
    CK(cudaMemset(b, 0, n * 4));
    float m = timeit([&] { cudaMemsetAsync(b, 0, n * 4); });
    float c = timeit([&] { cudaMemcpyAsync(b, a, n * 4, cudaMemcpyDeviceToDevice); });
    printf("cudaMemset %7.3f ms %7.1f GB/s | cudaMemcpy D2D %7.3f ms %7.1f GB/s\n", m, 4.0 * n / m / 1e6, c, 8.0 * n / c / 1e6);
    return 0;
}
