// scalix-b200: streaming single-pass inclusive scan for sm_100a.
//
// Same contract as scan.cuh (one kernel per device replacing the reference's three-stage
// scan, /root/reference scalix/algorithm/inclusive_scan.cuh:75-157) but built so that HBM
// never waits for the look-back:
//
//   * persistent CTAs (a multiple of the 148 SMs); tiles are handed out in scan order by an
//     atomic ticket, so a tile only ever waits for tiles that are already running;
//   * every tile is moved global -> shared memory by the TMA unit (cp.async.bulk, 1-D) into
//     a ring of STAGES buffers that is filled AHEAD of the consumer: while the CTA scans
//     tile i (and waits for its predecessors' prefixes), the loads of tiles i+1 .. are
//     already in flight and the store of tile i-1 is draining;
//   * the tile is scanned IN PLACE in shared memory (thread t owns VECS consecutive 16-byte
//     vectors; VECS odd => conflict-free LDS.128/STS.128) and written back with one bulk
//     store (cp.async.bulk.global.shared::cta), so the SM issues no per-thread global
//     loads/stores at all on full tiles;
//   * tile prefixes travel through packed {epoch|state, value} descriptors exactly like
//     scan.cuh (decoupled look-back, one relaxed 64/128-bit store/load per publication).
//
// Algorithmic bytes: sizeof(T) read + sizeof(T) written per element; nothing is re-read.
//
// Operand order matches scan.cuh / the reference: op(earlier, later) inside the device,
// carry applied as op(x, carry) except on the device's last element, op(carry, x)
// (/root/reference scalix/algorithm/inclusive_scan.cuh:122,150).
#pragma once

#include "scan.cuh"

namespace sclx::b200 {

// ---------------------------------------------------------------------------
// PTX wrappers: mbarrier + 1-D bulk copies
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
    return static_cast<unsigned>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void
bulk_load(void* smem_dst, const void* gmem_src, unsigned bytes, unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gmem_src),
        "r"(bytes),
        "r"(smem_u32(bar))
        : "memory");
}
// shared -> global, tracked by the thread's bulk async-group
__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template<int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (TMA)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---------------------------------------------------------------------------
// decoupled look-back (warp-wide; call with all 32 lanes of one warp, tile >= 1)
// returns op-fold of the inclusive totals of tiles [0, tile) in every lane
// ---------------------------------------------------------------------------
template<class T, class Op>
__device__ __forceinline__ T lookback_prefix(
    typename tile_desc<T>::word_t* desc,
    long long tile,
    unsigned st_aggregate,
    unsigned st_prefix,
    int lane,
    Op& op
) {
    using desc_io   = tile_desc<T>;
    long long first = tile - 1;
    bool have_acc   = false;
    T acc{};
    while (true) {
        const long long t = first - lane;
        unsigned st       = st_prefix;  // out-of-range lanes: never selected
        T val{};
        const bool in_range = t >= 0;
        if (in_range) {
            do {
                desc_io::load(desc + t, st, val);
            } while (st != st_aggregate && st != st_prefix);
        }
        const unsigned prefix_mask = __ballot_sync(kFullMask, in_range && st == st_prefix);
        // tile 0 always ends as PREFIX, so a window that reaches below tile 0 has a hit
        const int p = prefix_mask ? (__ffs(prefix_mask) - 1) : 31;
        // ordered fold of lanes p..0 (higher lane = earlier tile = left operand)
        T v = val;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            T y = shfl_down(v, d);
            if (lane + d <= p) v = op(y, v);
        }
        v        = shfl_idx(v, 0);
        acc      = have_acc ? op(v, acc) : v;
        have_acc = true;
        if (prefix_mask) break;
        first -= 32;
    }
    return acc;
}

// ---------------------------------------------------------------------------
// geometry
// ---------------------------------------------------------------------------
template<class T, int THREADS, int VECS, int STAGES>
struct scan_stream_config {
    static constexpr int V               = vec_traits<T>::width;
    static constexpr int warps           = THREADS / 32;
    static constexpr int items_per_thr   = V * VECS;
    static constexpr size_t tile_items   = size_t(THREADS) * items_per_thr;
    static constexpr size_t tile_bytes   = tile_items * sizeof(T);
    static constexpr size_t smem_bytes   = size_t(STAGES) * tile_bytes + 128;  // + alignment slack
    static_assert(sizeof(T) * V == 16, "tiles are made of 16-byte vectors");
    static_assert(VECS % 2 == 1, "VECS must be odd: thread-contiguous 16-byte vectors are then "
                                 "bank-conflict free");
    static_assert(THREADS % 32 == 0 && THREADS <= 1024, "whole warps");
    static_assert(STAGES >= 2, "need at least a double buffer");
};

struct scan_stream_counters {
    unsigned ticket;  // next tile to hand out
    unsigned exited;  // CTAs that have finished (the last one resets both)
};

// `in`/`out` must be 16-byte aligned (bulk copies); callers fall back to
// inclusive_scan_kernel<..., ALIGNED=false> otherwise.  `in` may equal `out`.
template<class T, class Op, int THREADS, int VECS, int STAGES>
__global__ void __launch_bounds__(THREADS) inclusive_scan_stream_kernel(
    const T* in,
    T* out,
    size_t n,
    typename tile_desc<T>::word_t* __restrict__ desc,
    unsigned epoch,
    scan_carry_t carry_in,
    T* __restrict__ total_out,
    scan_stream_counters* __restrict__ counters,
    Op op
) {
    using cfg     = scan_stream_config<T, THREADS, VECS, STAGES>;
    using desc_io = tile_desc<T>;
    constexpr int V     = cfg::V;
    constexpr int WARPS = cfg::warps;
    constexpr int IPT   = cfg::items_per_thr;

    extern __shared__ unsigned char s_dyn[];
    // stage buffers, 128-byte aligned
    unsigned char* s_tiles = reinterpret_cast<unsigned char*>(
        (reinterpret_cast<uintptr_t>(s_dyn) + 127) & ~uintptr_t(127));

    __shared__ __align__(8) unsigned long long s_full[STAGES];
    __shared__ long long s_tile_id[STAGES];
    __shared__ T s_warp_total[WARPS];
    __shared__ T s_warp_prefix[WARPS];

    const int tid  = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    const size_t num_tiles      = (n + cfg::tile_items - 1) / cfg::tile_items;
    const unsigned st_aggregate = (epoch << 2) | kTileAggregate;
    const unsigned st_prefix    = (epoch << 2) | kTilePrefix;

    // producer (thread 0): take the next ticket and start its load into `stage`
    auto produce = [&](int stage) {
        const size_t t = atomicAdd(&counters->ticket, 1u);
        if (t >= num_tiles) {
            s_tile_id[stage] = -1;
            mbar_arrive(&s_full[stage]);
            return;
        }
        s_tile_id[stage]    = (long long)t;
        const size_t base   = t * cfg::tile_items;
        const size_t valid  = (n - base < cfg::tile_items) ? n - base : cfg::tile_items;
        const size_t bytes  = valid * sizeof(T);
        const unsigned bulk = unsigned(bytes & ~size_t(15));
        T* dst              = reinterpret_cast<T*>(s_tiles + size_t(stage) * cfg::tile_bytes);
        // ragged end of the last tile (< 16 bytes): plain loads, published by the arrive below
        for (size_t i = bulk / sizeof(T); i < valid; ++i) dst[i] = in[base + i];
        if (bulk) {
            mbar_arrive_expect_tx(&s_full[stage], bulk);
            bulk_load(dst, in + base, bulk, &s_full[stage]);
        } else {
            mbar_arrive(&s_full[stage]);
        }
    };

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(&s_full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async_smem();
    }
    __syncthreads();
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES - 1; ++s) produce(s);
    }

    // cross-device carry, chained like the reference's host loop (ibid. :105-123)
    const bool has_carry = carry_in.count > 0;
    T carry{};
    if (has_carry) {
        const T* tot = static_cast<const T*>(carry_in.totals);
        carry        = tot[0];
        for (int i = 1; i < carry_in.count; ++i) carry = op(carry, tot[i]);
    }

    for (unsigned it = 0;; ++it) {
        const int stage = int(it % STAGES);
        mbar_wait(&s_full[stage], (it / STAGES) & 1u);
        const long long tile = s_tile_id[stage];
        if (tile < 0) break;

        const size_t tile_base = size_t(tile) * cfg::tile_items;
        const size_t valid = (n - tile_base < cfg::tile_items) ? n - tile_base : cfg::tile_items;
        T* s_tile = reinterpret_cast<T*>(s_tiles + size_t(stage) * cfg::tile_bytes);
        T* mine   = s_tile + size_t(tid) * IPT;

        // ---- shared -> registers, thread-serial scan -------------------------------
        T x[IPT];
#pragma unroll
        for (int k = 0; k < VECS; ++k) {
            const uint4 r = *reinterpret_cast<const uint4*>(mine + k * V);
            memcpy(&x[k * V], &r, 16);
        }
#pragma unroll
        for (int i = 1; i < IPT; ++i) x[i] = op(x[i - 1], x[i]);

        // ---- warp scan of thread totals -----------------------------------------------
        const T incl = warp_inclusive_scan(x[IPT - 1], op, lane);
        const T lex  = shfl_up(incl, 1);  // exclusive prefix inside the warp (lane > 0)
        if (lane == 31) s_warp_total[warp] = incl;
        __syncthreads();

        // ---- warp 0: block scan, publish, look back ----------------------------------------
        if (warp == 0) {
            const T wt          = s_warp_total[lane < WARPS ? lane : WARPS - 1];
            const T wi          = warp_inclusive_scan(wt, op, lane);
            const T block_total = shfl_idx(wi, WARPS - 1);
            const T wex         = shfl_up(wi, 1);  // exclusive per warp (lane > 0)
            T tile_prefix{};
            if (tile == 0) {
                if (lane == 0) desc_io::store(desc, st_prefix, block_total);
            } else {
                if (lane == 0) desc_io::store(desc + tile, st_aggregate, block_total);
                tile_prefix = lookback_prefix<T, Op>(desc, tile, st_aggregate, st_prefix, lane, op);
                if (lane == 0) desc_io::store(desc + tile, st_prefix, op(tile_prefix, block_total));
            }
            if (lane < WARPS)
                s_warp_prefix[lane]
                    = (lane == 0) ? tile_prefix : combine_left(tile > 0, tile_prefix, wex, op);
        }
        __syncthreads();

        // ---- apply prefixes, registers -> shared (in place) ---------------------------------
        {
            const bool warp_has = (tile > 0) || (warp > 0);
            bool has            = warp_has || lane > 0;
            T pre{};
            if (warp_has) {
                pre = s_warp_prefix[warp];
                if (lane > 0) pre = op(pre, lex);
            } else if (lane > 0) {
                pre = lex;
            }
            const size_t g0 = tile_base + size_t(tid) * IPT;
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                T r = has ? op(pre, x[i]) : x[i];
                if (has_carry) r = (g0 + i == n - 1) ? op(carry, r) : op(r, carry);
                x[i] = r;
            }
#pragma unroll
            for (int k = 0; k < VECS; ++k) {
                uint4 r;
                memcpy(&r, &x[k * V], 16);
                *reinterpret_cast<uint4*>(mine + k * V) = r;
            }
            const size_t bulk_items = (valid * sizeof(T) & ~size_t(15)) / sizeof(T);
            if (valid != cfg::tile_items) {
                // ragged end of the last tile: the < 16 bytes the bulk store cannot carry
#pragma unroll
                for (int i = 0; i < IPT; ++i) {
                    const size_t li = size_t(tid) * IPT + i;
                    if (li >= bulk_items && li < valid) out[tile_base + li] = x[i];
                }
            }
            if (total_out != nullptr) {
#pragma unroll
                for (int i = 0; i < IPT; ++i)
                    if (g0 + i == n - 1) *total_out = x[i];
            }
        }
        fence_proxy_async_smem();
        __syncthreads();

        // ---- producer: drain this tile, refill the stage drained one iteration ago ------------
        if (tid == 0) {
            const unsigned bulk = unsigned(valid * sizeof(T) & ~size_t(15));
            if (bulk) bulk_store(out + tile_base, s_tile, bulk);
            bulk_commit();
            bulk_wait_read<1>();  // the store issued in the previous iteration has left smem
            produce(int((it + STAGES - 1) % STAGES));
        }
    }

    if (tid == 0) {
        bulk_wait_all();
        __threadfence();
        const unsigned done = atomicAdd(&counters->exited, 1u);
        if (done == gridDim.x - 1) {
            // every CTA has taken its last ticket: ready for the next launch on this stream
            counters->ticket = 0;
            __threadfence();
            counters->exited = 0;
        }
    }
}

}  // namespace sclx::b200
