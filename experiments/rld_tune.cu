// Experiment: reduce_last_dim_kernel geometry for config 2 ([64, 2^24] f32): loads in flight
// (UNROLL), accumulator sets (ACCS -> registers -> resident CTAs) and CTAs per SM.
#include "scalix/detail/b200/reduce.cuh"
#include <cstdio>
#include <cstdlib>
using namespace sclx::b200;
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e__)); exit(2); } } while (0)
struct plus_op { __device__ float operator()(float a, float b) const { return a + b; } };
template<int UNROLL, int ACCS>
void run(const float* in, float* out, void* partials, unsigned* ticket) {
    const size_t rows = 64, cols = size_t(1) << 24;
    auto k = reduce_last_dim_kernel<float, plus_op, 256, UNROLL, 4, ACCS>;
    int fit = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fit, k, 256, 256 * 16);
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k);
    for (int per_sm : {2, 4, 6, 8, 12, 16}) {
        if (per_sm > fit * 2) continue;
        const unsigned grid = 148u * per_sm;
        reduce_workspace_t ws{partials, ticket};
        auto launch = [&] { k<<<grid, 256, 256 * 16>>>(in, rows, cols, 0.f, out, ws, plus_op{}); };
        cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
        launch(); launch(); CK(cudaDeviceSynchronize());
        cudaEventRecord(s); for (int r = 0; r < 10; ++r) launch(); cudaEventRecord(e); CK(cudaEventSynchronize(e));
        float ms; cudaEventElapsedTime(&ms, s, e); ms /= 10;
        printf("unroll %2d accs %d regs %3d fit %d  %2d CTA/SM  %7.4f ms  %7.1f GB/s\n", UNROLL, ACCS, fa.numRegs, fit, per_sm, ms,
               4.0 * rows * cols / ms / 1e6);
        fflush(stdout);
    }
}
int main() {
    float *in, *out; void* partials; unsigned* ticket;
    CK(cudaMalloc(&in, size_t(4) << 30)); CK(cudaMemset(in, 0, size_t(4) << 30));
    CK(cudaMalloc(&out, 4096)); CK(cudaMalloc(&partials, 64 << 20)); CK(cudaMalloc(&ticket, 256)); CK(cudaMemset(ticket, 0, 256));
    run<8, 8>(in, out, partials, ticket);
    run<8, 2>(in, out, partials, ticket);
    run<8, 1>(in, out, partials, ticket);
    run<4, 1>(in, out, partials, ticket);
    run<4, 4>(in, out, partials, ticket);
    run<16, 2>(in, out, partials, ticket);
    run<16, 1>(in, out, partials, ticket);
    return 0;
}
