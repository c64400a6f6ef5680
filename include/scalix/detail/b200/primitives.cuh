// scalix-b200: device-side building blocks shared by every framework-owned kernel.
//
// Everything here is sm_100a SIMT code for HBM-bound integer/float streaming work:
// 128-bit global accesses, warp-shuffle scans/reductions that never assume the
// operator has an identity (the reference's scan ignores its `identity` argument,
// /root/reference scalix/algorithm/inclusive_scan.cuh:48,87-93), and ordered
// combines so non-commutative operators keep the reference's operand order.
#pragma once
#include <cstdlib>

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace sclx::b200 {

constexpr unsigned kFullMask = 0xffffffffu;
constexpr int kSmCount       = 148;  // B200: 2 dies x 74 SMs

// ---------------------------------------------------------------------------
// 16-byte vectors of T
// ---------------------------------------------------------------------------
template<class T>
struct vec_traits {
    static constexpr int bytes = 16;
    static constexpr int width = (sizeof(T) <= 16 && 16 % sizeof(T) == 0)
                                   ? int(16 / sizeof(T))
                                   : 1;
};

template<class T, int V = vec_traits<T>::width>
struct alignas(sizeof(T) * V >= 16 ? 16 : sizeof(T) * V) vec_t {
    T v[V];
};

template<class T>
__host__ __device__ __forceinline__ bool is_aligned16(const T* p) {
    return (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

// streaming 128-bit load: read-once data, do not pollute L1
template<class T, int V>
__device__ __forceinline__ vec_t<T, V> load_vec_stream(const T* p) {
    static_assert(sizeof(T) * V == 16, "128-bit vectors only");
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    vec_t<T, V> out;
    *reinterpret_cast<uint4*>(&out) = r;
    return out;
}

template<class T, int V>
__device__ __forceinline__ void store_vec(T* p, const vec_t<T, V>& x) {
    static_assert(sizeof(T) * V == 16, "128-bit vectors only");
    const uint4 r = *reinterpret_cast<const uint4*>(&x);
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p),
                 "r"(r.x),
                 "r"(r.y),
                 "r"(r.z),
                 "r"(r.w)
                 : "memory");
}

// L2-coherent load (bypasses the non-coherent L1): used to read values another
// CTA published before a __threadfence + atomic ticket
template<class T>
__device__ __forceinline__ T load_cg(const T* p) {
    T v;
    if constexpr (sizeof(T) == 16 && alignof(T) >= 16) {
        uint4 b = __ldcg(reinterpret_cast<const uint4*>(p));
        memcpy(&v, &b, 16);
    } else if constexpr (sizeof(T) == 8 && alignof(T) >= 8) {
        unsigned long long b = __ldcg(reinterpret_cast<const unsigned long long*>(p));
        memcpy(&v, &b, 8);
    } else if constexpr (sizeof(T) == 4 && alignof(T) >= 4) {
        unsigned b = __ldcg(reinterpret_cast<const unsigned*>(p));
        memcpy(&v, &b, 4);
    } else if constexpr (sizeof(T) == 2 && alignof(T) >= 2) {
        unsigned short b = __ldcg(reinterpret_cast<const unsigned short*>(p));
        memcpy(&v, &b, 2);
    } else {
        const volatile unsigned char* src = reinterpret_cast<const volatile unsigned char*>(p);
        unsigned char* dst                = reinterpret_cast<unsigned char*>(&v);
        for (size_t i = 0; i < sizeof(T); ++i) dst[i] = src[i];
    }
    return v;
}

// ---------------------------------------------------------------------------
// shuffles for arbitrary trivially-copyable T (1..16 bytes)
// ---------------------------------------------------------------------------
template<class T>
__device__ __forceinline__ T shfl_up(const T& x, int delta) {
    constexpr int words = (sizeof(T) + 3) / 4;
    union {
        T t;
        unsigned w[words];
    } in{}, out{};
    in.t = x;
#pragma unroll
    for (int i = 0; i < words; ++i)
        out.w[i] = __shfl_up_sync(kFullMask, in.w[i], delta);
    return out.t;
}

template<class T>
__device__ __forceinline__ T shfl_down(const T& x, int delta) {
    constexpr int words = (sizeof(T) + 3) / 4;
    union {
        T t;
        unsigned w[words];
    } in{}, out{};
    in.t = x;
#pragma unroll
    for (int i = 0; i < words; ++i)
        out.w[i] = __shfl_down_sync(kFullMask, in.w[i], delta);
    return out.t;
}

template<class T>
__device__ __forceinline__ T shfl_idx(const T& x, int src) {
    constexpr int words = (sizeof(T) + 3) / 4;
    union {
        T t;
        unsigned w[words];
    } in{}, out{};
    in.t = x;
#pragma unroll
    for (int i = 0; i < words; ++i)
        out.w[i] = __shfl_sync(kFullMask, in.w[i], src);
    return out.t;
}

// inclusive scan over the 32 lanes; lane i ends with op(x0, ..., xi) in lane order.
// No identity needed: lanes only combine with real predecessors.
template<class T, class Op>
__device__ __forceinline__ T warp_inclusive_scan(T x, Op& op, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T y = shfl_up(x, d);
        if (lane >= d) x = op(y, x);
    }
    return x;
}

// unordered full-warp reduction (for commutative reductions); result in every lane
template<class T, class Op>
__device__ __forceinline__ T warp_reduce(T x, Op& op) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        T y = shfl_down(x, d);
        x   = op(x, y);
    }
    return shfl_idx(x, 0);
}

// combine an optional left prefix with a value: (has ? op(left, x) : x)
template<class T, class Op>
__device__ __forceinline__ T
combine_left(bool has_left, const T& left, const T& x, Op& op) {
    return has_left ? op(left, x) : x;
}

// Grid sizing: `waves` resident waves of ctas_per_sm CTAs per SM (a multiple of the SM count),
// capped by the work available.  One wave (a truly persistent grid) is NOT the fastest shape
// for the streaming kernels on B200: CTAs of later waves start staggered and de-synchronise the
// machine's load and store phases — fill 6037 -> 7518 GB/s, a copy-like generate 5157 -> 6955
// GB/s, axpy lambda 4747 -> 4942 GB/s from 1 to 64-256 waves (profiles/r01_fill_tune.txt,
// r01_dispatch_tune.txt).  SCLX_B200_WAVES overrides every kernel's default (tuning only).
// write-dominated streaming kernels (fill, iota, generate): effectively one CTA per chunk
constexpr unsigned kStreamWaves = 256;

__host__ inline unsigned tuning_waves(unsigned dflt) {
    static const int forced = [] {
        const char* e = std::getenv("SCLX_B200_WAVES");
        return e ? std::atoi(e) : 0;
    }();
    return forced > 0 ? unsigned(forced) : dflt;
}

__host__ inline unsigned
persistent_grid(size_t work_items, unsigned ctas_per_sm, int sm_count = kSmCount, unsigned waves = 1) {
    size_t cap = size_t(sm_count) * ctas_per_sm * tuning_waves(waves);
    if (cap > 0x7fffffffull) cap = 0x7fffffffull;
    if (work_items < 1) work_items = 1;
    return unsigned(work_items < cap ? work_items : cap);
}

}  // namespace sclx::b200
