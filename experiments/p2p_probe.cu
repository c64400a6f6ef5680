// Experiment: what does NVLink 5 P2P give for the access shapes a cross-partition gather can use?
// Two GPUs, GPU 0 runs the kernels, `remote` lives on GPU 1 (peer access enabled).
//   seq_read      coalesced 16 B loads of remote memory (bulk bandwidth)
//   seq_write     coalesced 16 B stores to remote memory
//   rand_read4/U  random 4 B loads, U independent loads in flight per thread
//   rand_read16   random 16 B loads
//   rand_line     each warp reads one random 128 B line (32 x 4 B)
//   rand_write4   random 4 B stores (posted)
//   rand_red4     random 4 B atomic adds (no return)
// Same kernels against a LOCAL buffer for comparison.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e__)); exit(2); } } while (0)

__device__ __forceinline__ unsigned long long mix(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void seq_read(const uint4* __restrict__ p, size_t n16, unsigned* sink) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    unsigned acc = 0;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        uint4 a = p[i], b = p[i + stride], c = p[i + 2 * stride], d = p[i + 3 * stride];
        acc += a.x ^ b.y ^ c.z ^ d.w;
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void seq_write(uint4* p, size_t n16) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    for (; i < n16; i += stride) p[i] = make_uint4(unsigned(i), 1, 2, 3);
}
template<int U>
__global__ void rand_read4(const unsigned* __restrict__ p, size_t n_elems, size_t n_req, unsigned* sink) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    unsigned acc = 0;
    for (; i < n_req; i += stride * U) {
        unsigned v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = p[mix(i + u * stride) % n_elems];
#pragma unroll
        for (int u = 0; u < U; ++u) acc += v[u];
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void rand_read16(const uint4* __restrict__ p, size_t n16, size_t n_req, unsigned* sink) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    unsigned acc = 0;
    for (; i < n_req; i += stride * 4) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = p[mix(i + u * stride) % n16];
#pragma unroll
        for (int u = 0; u < 4; ++u) acc += v[u].x ^ v[u].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void rand_line(const unsigned* __restrict__ p, size_t n_lines, size_t n_req_lines, unsigned* sink) {
    size_t w = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32, nw = size_t(gridDim.x) * blockDim.x / 32;
    unsigned lane = threadIdx.x & 31, acc = 0;
    for (; w < n_req_lines; w += nw * 4) {
        unsigned v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = p[(mix(w + u * nw) % n_lines) * 32 + lane];
#pragma unroll
        for (int u = 0; u < 4; ++u) acc += v[u];
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void rand_write4(unsigned* p, size_t n_elems, size_t n_req) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    for (; i < n_req; i += stride) p[mix(i) % n_elems] = unsigned(i);
}
__global__ void rand_red4(unsigned* p, size_t n_elems, size_t n_req) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
    for (; i < n_req; i += stride) atomicAdd(&p[mix(i) % n_elems], 1u);
}

template<class L>
void timeit(const char* name, const char* where, double useful_bytes, double reqs, L launch) {
    cudaEvent_t s, e; cudaEventCreate(&s); cudaEventCreate(&e);
    launch(); CK(cudaDeviceSynchronize());
    cudaEventRecord(s);
    const int reps = 3;
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e); CK(cudaEventSynchronize(e));
    float ms; cudaEventElapsedTime(&ms, s, e); ms /= reps;
    printf("%-14s %-6s %9.3f ms  %8.1f GB/s useful  %7.2f Greq/s\n", name, where, ms, useful_bytes / ms / 1e6, reqs / ms / 1e6);
    fflush(stdout);
}

int main() {
    int ndev = 0; CK(cudaGetDeviceCount(&ndev));
    const size_t bytes = size_t(2) << 30;  // 2 GiB target buffer
    unsigned *local = nullptr, *remote = nullptr, *sink = nullptr;
    CK(cudaSetDevice(0)); CK(cudaMalloc(&local, bytes)); CK(cudaMalloc(&sink, 4)); CK(cudaMemset(local, 0, bytes));
    if (ndev > 1) {
        CK(cudaSetDevice(1)); CK(cudaMalloc(&remote, bytes)); CK(cudaMemset(remote, 0, bytes));
        CK(cudaDeviceSynchronize());
        CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0));
    }
    const int grid = 148 * 8, block = 256;
    const size_t n_elems = bytes / 4, n16 = bytes / 16, n_lines = bytes / 128, n_req = size_t(1) << 26;
    for (int pass = 0; pass < 2; ++pass) {
        unsigned* p = pass == 0 ? local : remote;
        const char* where = pass == 0 ? "local" : "remote";
        if (!p) break;
        timeit("seq_read", where, double(bytes), double(bytes) / 128, [&] { seq_read<<<grid, block>>>((const uint4*)p, n16, sink); });
        timeit("seq_write", where, double(bytes), double(bytes) / 128, [&] { seq_write<<<grid, block>>>((uint4*)p, n16); });
        timeit("rand_read4/1", where, 4.0 * n_req, double(n_req), [&] { rand_read4<1><<<grid, block>>>(p, n_elems, n_req, sink); });
        timeit("rand_read4/4", where, 4.0 * n_req, double(n_req), [&] { rand_read4<4><<<grid, block>>>(p, n_elems, n_req, sink); });
        timeit("rand_read4/16", where, 4.0 * n_req, double(n_req), [&] { rand_read4<16><<<grid, block>>>(p, n_elems, n_req, sink); });
        timeit("rand_read16", where, 16.0 * n_req, double(n_req), [&] { rand_read16<<<grid, block>>>((const uint4*)p, n16, n_req, sink); });
        timeit("rand_line", where, 128.0 * (n_req / 8), double(n_req / 8), [&] { rand_line<<<grid, block>>>(p, n_lines, n_req / 8, sink); });
        timeit("rand_write4", where, 4.0 * n_req, double(n_req), [&] { rand_write4<<<grid, block>>>(p, n_elems, n_req); });
        timeit("rand_red4", where, 4.0 * n_req, double(n_req), [&] { rand_red4<<<grid, block>>>(p, n_elems, n_req); });
    }
    return 0;
}
