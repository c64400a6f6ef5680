// scalix-b200: framework-owned streaming kernels whose access pattern is known,
// so they use 128-bit accesses directly instead of going through the per-element
// lambda dispatcher.
//
//  * fill_kernel   arr[i] = value          (/root/reference scalix/fill.cuh:40-51)
//  * iota_kernel   arr[i] = T(start) + i   (/root/reference scalix/iota.cuh:40-51)
//  * gather_kernel out[i] = src[idx[i]]    (the access pattern of
//                  /root/reference examples/scaling_performance.cu:59-78 and
//                  examples/distributed_indexing.cu:146-154 with the arithmetic removed)
#pragma once

#include "primitives.cuh"

namespace sclx::b200 {

// Fill [out, out+n) with `value`.  Handles any alignment: scalar head up to the
// first 16-byte boundary, 128-bit body, scalar tail.
template<class T, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS)
    fill_kernel(T* __restrict__ out, size_t n, T value) {
    constexpr int V = vec_traits<T>::width;
    const size_t gtid = size_t(blockIdx.x) * THREADS + threadIdx.x;
    const size_t gstride = size_t(gridDim.x) * THREADS;

    size_t head = 0;
    if constexpr (V > 1) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(out);
        head              = ((16 - (a & 15u)) & 15u) / sizeof(T);
        if (head > n) head = n;
        const size_t nvec = (n - head) / V;
        vec_t<T, V> v;
#pragma unroll
        for (int e = 0; e < V; ++e) v.v[e] = value;
        T* body  = out + head;
        size_t i = gtid;
        for (; i + (UNROLL - 1) * gstride < nvec; i += UNROLL * gstride) {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
                store_vec<T, V>(body + (i + u * gstride) * V, v);
        }
        for (; i < nvec; i += gstride) store_vec<T, V>(body + i * V, v);
        // scalar head and tail
        if (gtid < head) out[gtid] = value;
        const size_t tail0 = head + nvec * V;
        if (tail0 + gtid < n) out[tail0 + gtid] = value;
    } else {
        for (size_t i = gtid; i < n; i += gstride) out[i] = value;
    }
}

// arr[i] = start + T(first_index + i): `first_index` is the partition's offset in
// the whole array so every device writes its own slice of one global sequence.
template<class T, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS)
    iota_kernel(T* __restrict__ out, size_t n, T start, size_t first_index) {
    constexpr int V = vec_traits<T>::width;
    const size_t gtid = size_t(blockIdx.x) * THREADS + threadIdx.x;
    const size_t gstride = size_t(gridDim.x) * THREADS;

    if constexpr (V > 1) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(out);
        size_t head       = ((16 - (a & 15u)) & 15u) / sizeof(T);
        if (head > n) head = n;
        const size_t nvec = (n - head) / V;
        T* body           = out + head;
        size_t i          = gtid;
        for (; i + (UNROLL - 1) * gstride < nvec; i += UNROLL * gstride) {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const size_t e0 = first_index + head + (i + u * gstride) * V;
                vec_t<T, V> v;
#pragma unroll
                for (int e = 0; e < V; ++e) v.v[e] = start + static_cast<T>(e0 + e);
                store_vec<T, V>(body + (i + u * gstride) * V, v);
            }
        }
        for (; i < nvec; i += gstride) {
            const size_t e0 = first_index + head + i * V;
            vec_t<T, V> v;
#pragma unroll
            for (int e = 0; e < V; ++e) v.v[e] = start + static_cast<T>(e0 + e);
            store_vec<T, V>(body + i * V, v);
        }
        if (gtid < head) out[gtid] = start + static_cast<T>(first_index + gtid);
        const size_t tail0 = head + nvec * V;
        if (tail0 + gtid < n)
            out[tail0 + gtid] = start + static_cast<T>(first_index + tail0 + gtid);
    } else {
        for (size_t i = gtid; i < n; i += gstride)
            out[i] = start + static_cast<T>(first_index + i);
    }
}

// Partition table for gathers whose source is split over several devices and
// NOT mapped into one contiguous virtual range (multi-process ranks that opened
// each other's slices through CUDA IPC).  Element e lives in
// part[e / elems_per_part] at offset e % elems_per_part.
struct gather_parts_t {
    const void* part[8];
    unsigned long long elems_per_part;  // ceil(last/ndev) * leading elements
    int nparts;                         // 1: plain pointer in part[0]
};

// one random element load with the L2::64B prefetch-size hint: the default fetches 126 B of
// DRAM per 4-byte random load, the hint 63 B (profiles/r01_fetch_granularity.txt)
template<class T>
__device__ __forceinline__ T load_random(const T* p) {
    if constexpr (sizeof(T) == 4) {
        unsigned v;
        asm volatile("ld.global.nc.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
        return __builtin_bit_cast(T, v);
    } else if constexpr (sizeof(T) == 8) {
        unsigned long long v;
        asm volatile("ld.global.nc.L2::64B.b64 %0, [%1];" : "=l"(v) : "l"(p));
        return __builtin_bit_cast(T, v);
    } else {
        return __ldg(p);
    }
}

// out[i] = src[idx[i]].  Each thread keeps UNROLL independent random loads in
// flight; index reads are coalesced.
template<class T, class I, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS) gather_kernel(
    const I* __restrict__ idx,
    size_t n,
    gather_parts_t src,
    T* __restrict__ out
) {
    // a CTA owns one CONTIGUOUS chunk of THREADS*UNROLL indices per trip (launched as one chunk
    // per CTA): index reads and result writes of a CTA stay in one DRAM page neighbourhood
    constexpr size_t CHUNK = size_t(THREADS) * UNROLL;
    for (size_t base = size_t(blockIdx.x) * CHUNK; base < n; base += size_t(gridDim.x) * CHUNK) {
        const size_t left = n - base;
        unsigned long long j[UNROLL];
        T v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const size_t i = size_t(u) * THREADS + threadIdx.x;
            j[u]           = i < left ? (unsigned long long)__ldg(idx + base + i) : 0ull;
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const size_t i = size_t(u) * THREADS + threadIdx.x;
            if (i < left) {
                if (src.nparts == 1) {
                    v[u] = load_random(static_cast<const T*>(src.part[0]) + j[u]);
                } else {
                    const unsigned long long p = j[u] / src.elems_per_part;
                    const unsigned long long o = j[u] - p * src.elems_per_part;
                    v[u] = load_random(static_cast<const T*>(src.part[p]) + o);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const size_t i = size_t(u) * THREADS + threadIdx.x;
            if (i < left) out[base + i] = v[u];
        }
    }
}

// out[i] = gen(first + i) for i in [0, n): the elementwise kernel behind transform,
// elementwise_reduce and assign_array.  `gen` only READS (it returns the value), so a thread
// gathers UNROLL results from coalesced, independent loads before it issues any store: the
// bytes in flight per thread are UNROLL times those of a per-element lambda whose store may
// alias its loads (the limit of the generic range dispatcher).
template<class R, class Gen, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS)
    generate_kernel(R* __restrict__ out, size_t n, size_t first, Gen gen) {
    const size_t tile = size_t(THREADS) * UNROLL;
    for (size_t base = size_t(blockIdx.x) * tile; base < n; base += size_t(gridDim.x) * tile) {
        R v[UNROLL];
        if (base + tile <= n) {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
                v[u] = gen(first + base + size_t(u) * THREADS + threadIdx.x);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) out[base + size_t(u) * THREADS + threadIdx.x] = v[u];
        } else {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const size_t i = base + size_t(u) * THREADS + threadIdx.x;
                if (i < n) v[u] = gen(first + i);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const size_t i = base + size_t(u) * THREADS + threadIdx.x;
                if (i < n) out[i] = v[u];
            }
        }
    }
}

}  // namespace sclx::b200
