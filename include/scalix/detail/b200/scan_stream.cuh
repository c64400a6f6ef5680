// scalix-b200: streaming single-pass inclusive scan for sm_100a.
//
// Same contract as scan.cuh (one kernel per device replacing the reference's three-stage
// scan, /root/reference scalix/algorithm/inclusive_scan.cuh:75-157) but organised as a
// warp-specialised pipeline so that neither HBM nor the output path ever waits for the
// decoupled look-back:
//
//   loader  (1 thread)  takes the next tile ticket (atomic, scan order) and starts a TMA bulk
//                       load (cp.async.bulk, 1-D) of the tile into a free stage of a ring of
//                       STAGES shared-memory buffers;
//   reducers (W_RED warps) fold every 32xVECS-vector chunk of a tile AS SOON AS IT LANDS; the
//                       warp that finishes last publishes the tile AGGREGATE at once (so no
//                       other tile's look-back ever waits for this CTA's own look-back);
//   look-back (W_LB warps, 1 is best) looks back (32*LB_K descriptors per L2 round trip),
//                       publishes the tile's inclusive PREFIX and leaves per-chunk prefixes
//                       in shared memory;
//   scanners (W_SCAN warps) scan each chunk in place (thread-serial + one warp scan), add the
//                       chunk prefix (and the cross-device carry) and write back to smem;
//   storer  (1 thread)  drains the finished stage with one bulk store
//                       (cp.async.bulk.global.shared::cta) and hands the stage back.
//
// All hand-offs are mbarriers; there is no CTA-wide barrier in the steady state.  Because
// aggregates AND prefixes are published when a tile ARRIVES (STAGES-1 tiles ahead of the
// point where its prefix is consumed), the look-back latency overlaps the scan/store of the
// tiles in front of it, and a tile only ever depends on tiles with smaller tickets whose
// loads were issued before its own => no dead-lock even if CTAs are not co-resident.
//
// Thread t of a chunk owns VECS consecutive 16-byte vectors; VECS odd makes the
// thread-contiguous LDS.128/STS.128 bank-conflict free.  Full tiles never touch global
// memory from the SM's load/store units: HBM traffic is sizeof(T) read + sizeof(T) written
// per element, moved entirely by the TMA unit.
//
// Operand order matches scan.cuh / the reference: op(earlier, later) inside the device,
// carry applied as op(x, carry) except on the device's last element, op(carry, x)
// (/root/reference scalix/algorithm/inclusive_scan.cuh:122,150).
#pragma once

#include "scan.cuh"

namespace sclx::b200 {

#ifdef SCLX_SCAN_STATS
// experiment-only instrumentation: [0] look-back cycles, [1] look-back rounds, [2] descriptor
// re-polls, [3] scanner wait cycles, [4] reducer wait cycles, [5] loader wait cycles,
// [6] storer wait cycles, [7] tiles
__device__ unsigned long long g_scan_stats[16];
#define SCLX_STAT_ADD(i, v) atomicAdd(&g_scan_stats[i], (unsigned long long)(v))
#define SCLX_STAT_CLOCK() clock64()
#else
#define SCLX_STAT_ADD(i, v) ((void)0)
#define SCLX_STAT_CLOCK() 0ll
#endif

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
// returns the op-fold of the block totals of tiles [0, tile) in every lane.
//
// One round inspects 32*K predecessors: lane l owns the K consecutive tiles
// first-l*K-(K-1) .. first-l*K, issues its K descriptor loads back to back (one L2
// round trip for the whole window) and folds them in order, stopping at the nearest
// tile that already carries an inclusive PREFIX.  With every resident CTA working on a
// different tile of the same "wave", the nearest prefix is typically about one grid
// (148..296 tiles) behind: K = 4..8 covers that in one or two round trips where a
// 32-wide window needs five to ten.
// ---------------------------------------------------------------------------
template<class T, class Op, int K>
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
        // newest tile of this lane's group; j walks towards older tiles
        const long long newest = first - (long long)lane * K;
        unsigned st[K];
        T val[K];
#pragma unroll
        for (int j = 0; j < K; ++j) {
            st[j] = st_prefix;
            if (newest - j >= 0) desc_io::load(desc + (newest - j), st[j], val[j]);
        }
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if (newest - j >= 0) {
                while (st[j] != st_aggregate && st[j] != st_prefix) {
                    desc_io::load(desc + (newest - j), st[j], val[j]);
                    SCLX_STAT_ADD(2, 1);
                }
            }
        }
        if (lane == 0) SCLX_STAT_ADD(1, 1);
        // ordered fold inside the lane: older tiles are LEFT operands; stop at a prefix
        bool lane_prefix = false;  // this lane's fold starts at an inclusive prefix
        bool lane_any    = false;  // this lane holds at least one in-range tile
        T v{};
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if (!lane_prefix && newest - j >= 0) {
                v           = lane_any ? op(val[j], v) : val[j];
                lane_any    = true;
                lane_prefix = (st[j] == st_prefix);
            }
        }
        // tile 0 always ends as PREFIX, so a window reaching below tile 0 has a hit
        const unsigned prefix_mask = __ballot_sync(kFullMask, lane_prefix);
        const int p                = prefix_mask ? (__ffs(prefix_mask) - 1) : 31;
        // ordered fold of lanes p..0 (higher lane = older tiles = left operand);
        // lanes <= p always hold in-range tiles
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            T y = shfl_down(v, d);
            if (lane + d <= p) v = op(y, v);
        }
        v        = shfl_idx(v, 0);
        acc      = have_acc ? op(v, acc) : v;
        have_acc = true;
        if (prefix_mask) break;
        first -= 32 * K;
    }
    return acc;
}


// ---------------------------------------------------------------------------
// geometry
// ---------------------------------------------------------------------------
template<class T, int VECS, int CHUNKS, int STAGES, int W_RED, int W_SCAN, int W_LB = 1>
struct scan_stream_config {
    static constexpr int V               = vec_traits<T>::width;
    static constexpr int items_per_thr   = V * VECS;
    static constexpr int chunk_items     = 32 * items_per_thr;
    static constexpr size_t tile_items   = size_t(CHUNKS) * chunk_items;
    static constexpr size_t tile_bytes   = tile_items * sizeof(T);
    static constexpr int warps           = 2 + W_LB + W_RED + W_SCAN;
    static constexpr int threads         = warps * 32;
    static constexpr size_t smem_bytes   = size_t(STAGES) * tile_bytes;
    static_assert(sizeof(T) * V == 16, "tiles are made of 16-byte vectors");
    static_assert(VECS % 2 == 1, "VECS must be odd: thread-contiguous 16-byte vectors are then "
                                 "bank-conflict free");
    static_assert(CHUNKS >= 1 && CHUNKS <= 32, "one look-back lane per chunk");
    static_assert(STAGES >= 2, "need at least a double buffer");
    static_assert(threads <= 1024, "too many warps");
};

// Production geometry per element width (profiles/r01_scan_stream_sweep{,2,3}.txt, 2^29
// elements on a B200): 36 KiB tiles of 8 chunks x 9 vectors per lane in a 6-stage ring
// (216 KiB), 4 reducer warps, 16 scanner warps, one look-back warp reading 32 descriptors per
// round trip -> 5941 GB/s for 4-byte and 6322 GB/s for 8-byte elements.  Narrower elements keep
// 5 vectors per lane so a lane's items (16/sizeof(T) per vector) stay in registers.
template<class T>
struct scan_stream_geometry {
    static constexpr bool wide   = sizeof(T) >= 4;
    static constexpr int vecs    = wide ? 9 : 5;
    static constexpr int chunks  = wide ? 8 : 16;
    static constexpr int stages  = wide ? 6 : 5;
    static constexpr int w_red   = wide ? 4 : 8;
    static constexpr int w_scan  = 16;
    static constexpr int lb_k    = 1;
    using config = scan_stream_config<T, vecs, chunks, stages, w_red, w_scan>;
};

struct scan_stream_counters {
    unsigned ticket;  // next tile to hand out
    unsigned exited;  // CTAs that have finished (the last one resets both)
};

// `in`/`out` must be 16-byte aligned (bulk copies); callers fall back to
// inclusive_scan_kernel<..., ALIGNED=false> otherwise.  `in` may equal `out`.
template<class T, class Op, int VECS, int CHUNKS, int STAGES, int W_RED, int W_SCAN, int LB_K,
         int W_LB = 1, int RUN = 1>
__global__ void __launch_bounds__((2 + W_LB + W_RED + W_SCAN) * 32) inclusive_scan_stream_kernel(
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
    using cfg     = scan_stream_config<T, VECS, CHUNKS, STAGES, W_RED, W_SCAN, W_LB>;
    using desc_io = tile_desc<T>;
    constexpr int V   = cfg::V;
    constexpr int IPT = cfg::items_per_thr;
    static_assert(RUN >= 1 && (RUN == 1 || W_LB == 1), "runs of tiles need the single look-back warp");
    static_assert(RUN <= STAGES, "a run must fit in the ring");

    extern __shared__ __align__(128) unsigned char s_tiles[];  // STAGES tile buffers

    __shared__ __align__(8) unsigned long long s_full[STAGES];      // TMA bytes landed
    __shared__ __align__(8) unsigned long long s_reduced[STAGES];   // chunk totals written
    __shared__ __align__(8) unsigned long long s_prefixed[STAGES];  // chunk prefixes written
    __shared__ __align__(8) unsigned long long s_written[STAGES];   // tile scanned in place
    __shared__ __align__(8) unsigned long long s_empty[STAGES];     // bulk store has left smem
    __shared__ long long s_tile_id[STAGES];
    __shared__ T s_chunk_total[STAGES][CHUNKS];   // per-chunk folds (reducers)
    __shared__ T s_chunk_incl[STAGES][CHUNKS];    // inclusive scan of the chunk folds
    __shared__ T s_chunk_prefix[STAGES][CHUNKS];  // exclusive prefix of each chunk, tile prefix folded in
    __shared__ unsigned s_red_count[STAGES];      // reducer warps done with the stage
#ifdef SCLX_SCAN_STATS
    __shared__ long long s_ts[STAGES];
#define SCLX_STAT_SINCE(i, stage) SCLX_STAT_ADD(i, clock64() - s_ts[stage])
#else
#define SCLX_STAT_SINCE(i, stage) ((void)0)
#endif

    const int tid  = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    const size_t num_tiles      = (n + cfg::tile_items - 1) / cfg::tile_items;
    const unsigned st_aggregate = (epoch << 2) | kTileAggregate;
    const unsigned st_prefix    = (epoch << 2) | kTilePrefix;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&s_full[s], 1);
            mbar_init(&s_reduced[s], 1);
            s_red_count[s] = 0;
            mbar_init(&s_prefixed[s], 1);
            mbar_init(&s_written[s], W_SCAN);
            mbar_init(&s_empty[s], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async_smem();
    }
    __syncthreads();

    auto stage_ptr = [&](int stage) {
        return reinterpret_cast<T*>(s_tiles + size_t(stage) * cfg::tile_bytes);
    };
    auto valid_items = [&](long long tile) {
        const size_t base = size_t(tile) * cfg::tile_items;
        return (n - base < cfg::tile_items) ? n - base : cfg::tile_items;
    };

    if (warp == 0) {
        // ================= loader ======================================================
        if (lane == 0) {
            int markers = 0;
            // a ticket is a RUN of consecutive tiles: the look-back warp derives the prefix of
            // the 2nd..RUNth tile from its own previous tile instead of looking back again.
            // Measured slower than RUN = 1 on B200 (profiles/r01_scan_stream_sweep{6,7}_*.txt);
            // production uses 1, the parameter stays for experiments/scan_tune2.cu.
            size_t run_base = 0;
            int run_left    = 0;
            for (unsigned it = 0;; ++it) {
                const int stage      = int(it % STAGES);
                const unsigned round = it / STAGES;
                [[maybe_unused]] const long long w0 = SCLX_STAT_CLOCK();
                if (RUN > 1 && markers == 0 && run_left == 0) {
                    // take a run only when all of its stages are free (stages recycle in order):
                    // a ticketed tile whose load has not been issued stalls every later tile's
                    // look-back
                    const unsigned last = it + unsigned(RUN) - 1u;
                    mbar_wait(&s_empty[last % STAGES], ((last / STAGES) & 1u) ^ 1u);
                }
                mbar_wait(&s_empty[stage], (round & 1u) ^ 1u);  // fresh barrier: passes
                SCLX_STAT_ADD(5, SCLX_STAT_CLOCK() - w0);
                size_t t = num_tiles;
                if (markers == 0) {
                    if (run_left == 0) {
                        run_base = atomicAdd(&counters->ticket, unsigned(RUN));
                        run_left = RUN;
                    }
                    t = run_base + size_t(RUN - run_left);
                    --run_left;
                }
                if (t >= num_tiles) {
                    // end of input: one marker per look-back warp flows through the pipeline
                    s_tile_id[stage] = -1;
                    mbar_arrive(&s_full[stage]);
                    if (++markers == W_LB) break;
                    continue;
                }
                s_tile_id[stage]    = (long long)t;
#ifdef SCLX_SCAN_STATS
                s_ts[stage] = clock64();
#endif
                const size_t base   = t * cfg::tile_items;
                const size_t valid  = valid_items((long long)t);
                const unsigned bulk = unsigned(valid * sizeof(T) & ~size_t(15));
                T* dst              = stage_ptr(stage);
                // ragged end of the last tile (< 16 bytes): plain loads, released by the arrive
                for (size_t i = bulk / sizeof(T); i < valid; ++i) dst[i] = in[base + i];
                if (bulk) {
                    mbar_arrive_expect_tx(&s_full[stage], bulk);
                    bulk_load(dst, in + base, bulk, &s_full[stage]);
                } else {
                    mbar_arrive(&s_full[stage]);
                }
            }
            __threadfence();
            const unsigned done = atomicAdd(&counters->exited, 1u);
            if (done == gridDim.x - 1) {
                // every loader has taken its last ticket: ready for the next launch
                counters->ticket = 0;
                __threadfence();
                counters->exited = 0;
            }
        }
    } else if (warp == 1) {
        // ================= storer ======================================================
        if (lane == 0) {
            int markers = 0;
            unsigned it = 0;
            for (;; ++it) {
                const int stage      = int(it % STAGES);
                const unsigned round = it / STAGES;
                [[maybe_unused]] const long long w0 = SCLX_STAT_CLOCK();
                mbar_wait(&s_written[stage], round & 1u);
                SCLX_STAT_ADD(6, SCLX_STAT_CLOCK() - w0);
                const long long tile = s_tile_id[stage];
                if (tile < 0) {
                    if (++markers == W_LB) break;
                    mbar_arrive(&s_empty[stage]);
                    continue;
                }
                SCLX_STAT_SINCE(12, stage);
                const unsigned bulk = unsigned(valid_items(tile) * sizeof(T) & ~size_t(15));
                if (bulk) bulk_store(out + size_t(tile) * cfg::tile_items, stage_ptr(stage), bulk);
                bulk_commit();
                bulk_wait_read<0>();  // the store has left shared memory: the stage is free
                SCLX_STAT_SINCE(13, stage);
                mbar_arrive(&s_empty[stage]);
            }
            bulk_wait_all();
        }
    } else if (warp < 2 + W_LB) {
        // ================= look-back (warp j takes every W_LB-th tile of this CTA) ======
        long long prev_tile = -2;  // previous tile of this CTA and its inclusive prefix
        T prev_inclusive{};
        for (unsigned it = unsigned(warp - 2);; it += W_LB) {
            const int stage      = int(it % STAGES);
            const unsigned round = it / STAGES;
            mbar_wait(&s_reduced[stage], round & 1u);
            const long long tile = s_tile_id[stage];
            if (tile < 0) {
                if (lane == 0) mbar_arrive(&s_prefixed[stage]);
                break;
            }
            // the aggregate (tile 0: the prefix) was already published by the reducers
            if (lane == 0) SCLX_STAT_SINCE(10, stage);
            const T own_total = s_chunk_incl[stage][CHUNKS - 1];
            T tile_prefix{};
            if (tile > 0) {
                [[maybe_unused]] const long long c0 = SCLX_STAT_CLOCK();
                if (RUN > 1 && tile == prev_tile + 1)
                    tile_prefix = prev_inclusive;  // same run: the predecessor was ours
                else
                    tile_prefix = lookback_prefix<T, Op, LB_K>(desc, tile, st_aggregate, st_prefix,
                                                               lane, op);
                if (lane == 0) {
                    SCLX_STAT_ADD(0, SCLX_STAT_CLOCK() - c0);
                    SCLX_STAT_ADD(7, 1);
                    desc_io::store(desc + tile, st_prefix, op(tile_prefix, own_total));
                }
            }
            prev_tile      = tile;
            prev_inclusive = tile > 0 ? op(tile_prefix, own_total) : own_total;
            const T cex = s_chunk_incl[stage][(lane > 0 && lane < CHUNKS) ? lane - 1 : 0];
            if (lane < CHUNKS)
                s_chunk_prefix[stage][lane]
                    = (lane == 0) ? tile_prefix : combine_left(tile > 0, tile_prefix, cex, op);
            __syncwarp();
            if (lane == 0) SCLX_STAT_SINCE(11, stage);
            if (lane == 0) mbar_arrive(&s_prefixed[stage]);
        }
    } else if (warp < 2 + W_LB + W_RED) {
        // ================= reducers ====================================================
        const int rw = warp - 2 - W_LB;
        int markers  = 0;
        for (unsigned it = 0;; ++it) {
            const int stage      = int(it % STAGES);
            const unsigned round = it / STAGES;
            [[maybe_unused]] const long long w0 = SCLX_STAT_CLOCK();
            mbar_wait(&s_full[stage], round & 1u);
            if (lane == 0 && rw == 0) SCLX_STAT_ADD(4, SCLX_STAT_CLOCK() - w0);
            const long long tile = s_tile_id[stage];
            if (tile < 0) {
                if (lane == 0) {
                    const unsigned arrived = atomicAdd(&s_red_count[stage], 1u);
                    if (arrived == W_RED - 1) {
                        s_red_count[stage] = 0;
                        mbar_arrive(&s_reduced[stage]);
                    }
                }
                if (++markers == W_LB) break;
                continue;
            }
            if (lane == 0 && rw == 0) SCLX_STAT_SINCE(8, stage);
            const T* s_tile = stage_ptr(stage);
            for (int c = rw; c < CHUNKS; c += W_RED) {
                const T* mine = s_tile + (size_t(c) * 32 + lane) * IPT;
                T x[IPT];
#pragma unroll
                for (int k = 0; k < VECS; ++k) {
                    const uint4 r = *reinterpret_cast<const uint4*>(mine + k * V);
                    memcpy(&x[k * V], &r, 16);
                }
                T a = x[0];
#pragma unroll
                for (int i = 1; i < IPT; ++i) a = op(a, x[i]);
                // ordered tree: lane 0 ends with op(a_0, ..., a_31)
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const T y = shfl_down(a, d);
                    if (lane + d < 32) a = op(a, y);
                }
                if (lane == 0) s_chunk_total[stage][c] = a;
            }
            // the reducer warp that finishes LAST turns the chunk folds into the tile aggregate
            // and publishes it at once: other tiles' look-backs never wait for this CTA's own
            // look-back to get round to this tile
            unsigned arrived = 0;
            __syncwarp();
            if (lane == 0) {
                __threadfence_block();
                arrived = atomicAdd(&s_red_count[stage], 1u);
            }
            arrived = __shfl_sync(kFullMask, arrived, 0);
            if (arrived == W_RED - 1) {
                __threadfence_block();
                const T ct = s_chunk_total[stage][lane < CHUNKS ? lane : CHUNKS - 1];
                const T ci = warp_inclusive_scan(ct, op, lane);
                if (lane < CHUNKS) s_chunk_incl[stage][lane] = ci;
                if (lane == CHUNKS - 1)
                    desc_io::store(desc + tile, tile == 0 ? st_prefix : st_aggregate, ci);
                if (lane == 0) s_red_count[stage] = 0;
                if (lane == 0) SCLX_STAT_SINCE(9, stage);
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_reduced[stage]);
            }
        }
    } else {
        // ================= scanners ====================================================
        const int sw = warp - 2 - W_LB - W_RED;
        // cross-device carry, chained like the reference's host loop (ibid. :105-123)
        const bool has_carry = carry_in.count > 0;
        T carry{};
        if (has_carry) {
            const T* tot = static_cast<const T*>(carry_in.totals);
            carry        = tot[0];
            for (int i = 1; i < carry_in.count; ++i) carry = op(carry, tot[i]);
        }
        int markers = 0;
        for (unsigned it = 0;; ++it) {
            const int stage      = int(it % STAGES);
            const unsigned round = it / STAGES;
            [[maybe_unused]] const long long w0 = SCLX_STAT_CLOCK();
            mbar_wait(&s_prefixed[stage], round & 1u);
            if (lane == 0 && sw == 0) SCLX_STAT_ADD(3, SCLX_STAT_CLOCK() - w0);
            const long long tile = s_tile_id[stage];
            if (tile < 0) {
                if (lane == 0) mbar_arrive(&s_written[stage]);
                if (++markers == W_LB) break;
                continue;
            }
            const size_t tile_base  = size_t(tile) * cfg::tile_items;
            const size_t valid      = valid_items(tile);
            const size_t bulk_items = (valid * sizeof(T) & ~size_t(15)) / sizeof(T);
            T* s_tile               = stage_ptr(stage);
            for (int c = sw; c < CHUNKS; c += W_SCAN) {
                const size_t local0 = (size_t(c) * 32 + lane) * IPT;
                T* mine             = s_tile + local0;
                T x[IPT];
#pragma unroll
                for (int k = 0; k < VECS; ++k) {
                    const uint4 r = *reinterpret_cast<const uint4*>(mine + k * V);
                    memcpy(&x[k * V], &r, 16);
                }
#pragma unroll
                for (int i = 1; i < IPT; ++i) x[i] = op(x[i - 1], x[i]);
                const T incl = warp_inclusive_scan(x[IPT - 1], op, lane);
                const T lex  = shfl_up(incl, 1);  // exclusive prefix inside the chunk (lane > 0)

                const bool chunk_has = (tile > 0) || (c > 0);
                const bool has       = chunk_has || lane > 0;
                T pre{};
                if (chunk_has) {
                    pre = s_chunk_prefix[stage][c];
                    if (lane > 0) pre = op(pre, lex);
                } else if (lane > 0) {
                    pre = lex;
                }
                const size_t g0 = tile_base + local0;
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
                if (valid != cfg::tile_items) {
                    // ragged end of the last tile: the < 16 bytes a bulk store cannot carry
#pragma unroll
                    for (int i = 0; i < IPT; ++i) {
                        const size_t li = local0 + i;
                        if (li >= bulk_items && li < valid) out[tile_base + li] = x[i];
                    }
                }
                if (total_out != nullptr && g0 + IPT >= n && g0 < n) {
#pragma unroll
                    for (int i = 0; i < IPT; ++i)
                        if (g0 + i == n - 1) *total_out = x[i];
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_written[stage]);
        }
    }
}

}  // namespace sclx::b200
