// scalix-b200: single-pass inclusive scan (decoupled look-back) for sm_100a.
//
// Replaces the reference's three-stage scan
//   /root/reference scalix/algorithm/inclusive_scan.cuh:75-103  (thrust::inclusive_scan per device)
//   /root/reference scalix/algorithm/inclusive_scan.cuh:105-123 (host carry chain through UM)
//   /root/reference scalix/algorithm/inclusive_scan.cuh:125-157 (second full fix-up pass)
// with ONE kernel per device that reads every element once and writes it once
// (algorithmic bytes: sizeof(Tin)+sizeof(Tout) per element), taking the
// cross-device carry as a device-side value so no host round trip is needed.
//
// Layout of one tile (THREADS x VECS 128-bit vectors):
//   warp w owns a contiguous run of 32*VECS vectors; in step k lane l holds vector
//   k*32+l of that run, so every warp-wide access is one fully coalesced 512-byte
//   request and the scan order is (warp, step, lane, element-in-vector).
// Tile prefixes travel through a descriptor array in global memory, one 64-bit
// (sizeof(T)<=4) or 128-bit (sizeof(T)==8) word per tile holding {epoch|state, value}
// written with a single relaxed store, so a reader can never see a torn pair.
// The epoch makes descriptors from earlier launches read as "invalid": the
// workspace is zeroed once at allocation and never again.
//
// Operand order (non-commutative operators): in-device combination is always
// op(earlier, later).  The cross-device carry follows the reference exactly:
// every element but the device's last becomes op(x, carry)
// (/root/reference scalix/algorithm/inclusive_scan.cuh:150) and the last one
// op(carry, x) (ibid. :122).
#pragma once

#include "primitives.cuh"

namespace sclx::b200 {

enum : unsigned { kTileInvalid = 0u, kTileAggregate = 1u, kTilePrefix = 2u };

// ---------------------------------------------------------------------------
// tile descriptors
// ---------------------------------------------------------------------------
template<class T, bool Small = (sizeof(T) <= 4)>
struct tile_desc;

template<class T>
struct tile_desc<T, true> {
    using word_t = unsigned long long;
    static constexpr size_t bytes = 8;

    static __device__ __forceinline__ void
    store(word_t* p, unsigned status, const T& value) {
        unsigned bits = 0;
        memcpy(&bits, &value, sizeof(T));
        word_t w = (word_t(status) << 32) | word_t(bits);
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w)
                     : "memory");
    }
    static __device__ __forceinline__ void
    load(const word_t* p, unsigned& status, T& value) {
        word_t w;
#ifdef SCLX_SCAN_WEAK_POLL
        asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
#else
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];"
                     : "=l"(w)
                     : "l"(p)
                     : "memory");
#endif
        status        = unsigned(w >> 32);
        unsigned bits = unsigned(w);
        memcpy(&value, &bits, sizeof(T));
    }
};

template<class T>
struct tile_desc<T, false> {
    static_assert(sizeof(T) == 8, "packed descriptors hold values of <= 8 bytes");
    struct alignas(16) word_t {
        unsigned long long s, v;
    };
    static constexpr size_t bytes = 16;

    static __device__ __forceinline__ void
    store(word_t* p, unsigned status, const T& value) {
        unsigned long long bits;
        memcpy(&bits, &value, 8);
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p),
                     "l"((unsigned long long)status),
                     "l"(bits)
                     : "memory");
    }
    static __device__ __forceinline__ void
    load(const word_t* p, unsigned& status, T& value) {
        unsigned long long s, bits;
#ifdef SCLX_SCAN_WEAK_POLL
        asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(s), "=l"(bits) : "l"(p) : "memory");
#else
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];"
                     : "=l"(s), "=l"(bits)
                     : "l"(p)
                     : "memory");
#endif
        status = unsigned(s);
        memcpy(&value, &bits, 8);
    }
};

template<class T>
constexpr size_t scan_desc_bytes() {
    return tile_desc<T>::bytes;
}

// ---------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------
template<class T, int THREADS, int VECS>
struct scan_config {
    static constexpr int V             = vec_traits<T>::width;
    static constexpr int warps         = THREADS / 32;
    static constexpr int items_per_thr = V * VECS;
    static constexpr size_t tile_items = size_t(THREADS) * items_per_thr;
};

struct scan_carry_t {
    // carry-in for partitions >= 1: `count` values in `totals` (device memory,
    // one per lower partition, in partition order) are chained as
    // c = op(c, totals[i]) exactly like the reference's host loop; count == 0
    // means "no carry" (partition 0 or single device).
    const void* totals;
    int count;
};

template<class T, class Op, int THREADS, int VECS, bool ALIGNED>
__global__ void __launch_bounds__(THREADS) inclusive_scan_kernel(
    const T* in,  // may alias `out` (in-place scan): no __restrict__
    T* out,
    size_t n,
    typename tile_desc<T>::word_t* __restrict__ desc,
    unsigned epoch,
    scan_carry_t carry_in,
    T* __restrict__ total_out,  // optional: this partition's (carried) last value
    Op op
) {
    using cfg  = scan_config<T, THREADS, VECS>;
    using desc_io = tile_desc<T>;
    constexpr int V     = cfg::V;
    constexpr int WARPS = cfg::warps;

    __shared__ T s_warp_total[WARPS];
    __shared__ T s_warp_prefix[WARPS];  // exclusive prefix of each warp (tile prefix folded in)
    __shared__ bool s_tile_has_prefix;

    const int tid  = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    const size_t tile      = blockIdx.x;
    const size_t tile_base = tile * cfg::tile_items;
    const size_t warp_base = tile_base + size_t(warp) * (32 * cfg::items_per_thr);
    const bool full_tile   = tile_base + cfg::tile_items <= n;

    const unsigned st_aggregate = (epoch << 2) | kTileAggregate;
    const unsigned st_prefix    = (epoch << 2) | kTilePrefix;

    // ---- load -------------------------------------------------------------
    T x[VECS][V];
    if (full_tile) {
#pragma unroll
        for (int k = 0; k < VECS; ++k) {
            const T* p = in + warp_base + (size_t(k) * 32 + lane) * V;
            if constexpr (ALIGNED && V > 1) {
                vec_t<T, V> v = load_vec_stream<T, V>(p);
#pragma unroll
                for (int e = 0; e < V; ++e) x[k][e] = v.v[e];
            } else {
#pragma unroll
                for (int e = 0; e < V; ++e) x[k][e] = p[e];
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < VECS; ++k) {
            const size_t g = warp_base + (size_t(k) * 32 + lane) * V;
#pragma unroll
            for (int e = 0; e < V; ++e) {
                // out-of-range slots replicate the last valid element read by
                // this thread (or stay unused); they are never stored and never
                // feed a valid element because they sit at the end of scan order
                if (g + e < n) x[k][e] = in[g + e];
                else x[k][e] = T{};
            }
        }
    }

    // ---- thread-local scan inside each vector ------------------------------
#pragma unroll
    for (int k = 0; k < VECS; ++k) {
#pragma unroll
        for (int e = 1; e < V; ++e) x[k][e] = op(x[k][e - 1], x[k][e]);
    }

    // ---- warp scans of the vector totals (VECS independent chains) ----------
    T incl[VECS];
#pragma unroll
    for (int k = 0; k < VECS; ++k)
        incl[k] = warp_inclusive_scan(x[k][V - 1], op, lane);

    // exclusive prefix of this thread's vector k inside the warp's run:
    // (steps 0..k-1 complete) ++ (lanes 0..lane-1 of step k)
    T tpre[VECS];  // valid unless (k == 0 && lane == 0)
    T run{};       // fold of completed steps
#pragma unroll
    for (int k = 0; k < VECS; ++k) {
        T up = shfl_up(incl[k], 1);
        if (k == 0) {
            tpre[k] = up;  // meaningless on lane 0 (flagged below)
        } else {
            tpre[k] = (lane == 0) ? run : op(run, up);
        }
        T step_total = shfl_idx(incl[k], 31);
        run          = (k == 0) ? step_total : op(run, step_total);
    }
    // `run` is now the warp's total
    if (lane == 31) s_warp_total[warp] = run;
    __syncthreads();

    // ---- warp 0: block scan of warp totals + decoupled look-back ------------
    if (warp == 0) {
        T wt = s_warp_total[lane < WARPS ? lane : WARPS - 1];
        T wi = warp_inclusive_scan(wt, op, lane);
        T block_total = shfl_idx(wi, WARPS - 1);
        T wex         = shfl_up(wi, 1);  // exclusive per warp (lane 0: none)

        bool has_tile_prefix = false;
        T tile_prefix{};
        if (tile == 0) {
            if (lane == 0) desc_io::store(desc, st_prefix, block_total);
        } else {
            if (lane == 0) desc_io::store(desc + tile, st_aggregate, block_total);

            // window of 32 predecessors, lane l looks at tile (first - l)
            long long first = (long long)tile - 1;
            bool have_acc   = false;
            T acc{};
            while (true) {
                const long long t = first - lane;
                unsigned st       = st_prefix;  // out-of-range lanes: never selected
                T val{};
                bool in_range = t >= 0;
                if (in_range) {
                    do {
                        desc_io::load(desc + t, st, val);
                    } while (st != st_aggregate && st != st_prefix);
                }
                // every lane now holds a valid descriptor of this epoch
                const unsigned prefix_mask
                    = __ballot_sync(kFullMask, in_range && st == st_prefix);
                // tile 0 always ends as PREFIX, so when the window reaches below
                // tile 0 prefix_mask is non-zero
                const int p = prefix_mask ? (__ffs(prefix_mask) - 1) : 31;
                // ordered fold of lanes p..0 (higher lane = earlier tile = left)
                T v = val;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    T y = shfl_down(v, d);
                    if (lane + d <= p) v = op(y, v);
                }
                v = shfl_idx(v, 0);
                // this window is earlier than everything accumulated so far
                acc      = have_acc ? op(v, acc) : v;
                have_acc = true;
                if (prefix_mask) break;
                first -= 32;
            }
            tile_prefix     = acc;
            has_tile_prefix = true;
            if (lane == 0)
                desc_io::store(desc + tile, st_prefix, op(tile_prefix, block_total));
        }

        if (lane < WARPS) {
            // exclusive prefix for warp `lane`
            T wp = (lane == 0) ? tile_prefix
                               : combine_left(has_tile_prefix, tile_prefix, wex, op);
            s_warp_prefix[lane] = wp;
        }
        if (lane == 0) s_tile_has_prefix = has_tile_prefix;
    }
    __syncthreads();

    // ---- apply prefixes ------------------------------------------------------
    const bool tile_has   = s_tile_has_prefix;
    const bool warp_has   = tile_has || warp > 0;
    const T warp_prefix   = s_warp_prefix[warp];

    // cross-device carry, chained like the reference's host loop
    bool has_carry = carry_in.count > 0;
    T carry{};
    if (has_carry) {
        const T* tot = static_cast<const T*>(carry_in.totals);
        carry        = tot[0];
        for (int i = 1; i < carry_in.count; ++i) carry = op(carry, tot[i]);
    }

#pragma unroll
    for (int k = 0; k < VECS; ++k) {
        const bool thr_has = (k > 0) || (lane > 0);
        T pre{};
        bool has = false;
        if (warp_has && thr_has) {
            pre = op(warp_prefix, tpre[k]);
            has = true;
        } else if (warp_has) {
            pre = warp_prefix;
            has = true;
        } else if (thr_has) {
            pre = tpre[k];
            has = true;
        }
        const size_t g = warp_base + (size_t(k) * 32 + lane) * V;
#pragma unroll
        for (int e = 0; e < V; ++e) {
            T r = has ? op(pre, x[k][e]) : x[k][e];
            if (has_carry) r = (g + e == n - 1) ? op(carry, r) : op(r, carry);
            x[k][e] = r;
        }
        if (full_tile) {
            T* p = out + g;
            if constexpr (ALIGNED && V > 1) {
                vec_t<T, V> v;
#pragma unroll
                for (int e = 0; e < V; ++e) v.v[e] = x[k][e];
                store_vec<T, V>(p, v);
            } else {
#pragma unroll
                for (int e = 0; e < V; ++e) p[e] = x[k][e];
            }
        } else {
#pragma unroll
            for (int e = 0; e < V; ++e)
                if (g + e < n) out[g + e] = x[k][e];
        }
        if (total_out != nullptr) {
#pragma unroll
            for (int e = 0; e < V; ++e)
                if (g + e == n - 1) *total_out = x[k][e];
        }
    }
}

}  // namespace sclx::b200
