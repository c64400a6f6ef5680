// scalix-b200: binned (exchange) form of the gather  out[i] = src[index[i]].
//
// Two measured limits shape it (experiments/p2p_probe.cu, experiments/fetch_granularity.cu;
// profiles/r01_p2p_probe.txt, r01_fetch_granularity.txt):
//   * NVLink 5 P2P serves ~6.5 G REQUESTS/s per GPU whatever a request carries, so one peer load
//     per index caps a cross-partition gather at ~26 GB/s of payload (3 % of the link);
//   * local HBM serves ~40 G random 4-byte reads/s whatever the fetch granularity (a DRAM
//     row-activation limit), although BASELINE config 4 wants one element per 16 source bytes,
//     i.e. every 32-byte sector about twice.
// Both go away when requests are BINNED by where their source element lives.  A bucket is one
// L2-sized REGION of one owner's partition (regions x owners <= kGxMaxBuckets):
//
//   K1  partition (requester r, one CTA per 4096-index tile): classify every index by bucket,
//       bucket the tile in shared memory, reserve a run in the (r -> bucket) request queue with
//       ONE atomic per bucket and write the 32-bit in-partition offsets to the queue IN THE
//       OWNER'S MEMORY with coalesced posted stores; keep the 16-bit in-tile positions locally,
//       in the same order, plus (first slot, length) of every run.
//   K2  serve (owner q), region after region: while region s is served, its CTAs pull region
//       s+1 into L2 with bulk prefetches (a sequential HBM stream); the gather loads of a region
//       then HIT L2 instead of opening a DRAM row each.  Values go, in queue order, to the
//       response buffer IN THE REQUESTER'S MEMORY (coalesced posted stores).
//   K3  place (requester r): per tile, stream the responses of its runs + positions and scatter
//       inside the tile's 16 KiB window of `out`.
//
// On one GPU this is a three-pass streaming gather (48 algorithmic bytes per index instead of a
// DRAM row activation); across GPUs only full-line NVLink writes cross the switch.  The three
// kernels are three launch groups ordered by stream events, no host round trip.  Queues have
// bounded capacity (1.25 x the uniform share + slack); an index that finds its queue full is
// served by a direct (peer) load in K1, so skewed index sets stay correct and merely lose the
// speed-up.  Pure data movement: results are bit-identical to the direct gather.
#pragma once

#include "primitives.cuh"

#include <cstdint>
#include <cstdio>

// -DSCLX_GX_CHECKS: bounds checks on every queue / response / shared-memory index of the three
// kernels; a violation traps the kernel (the launch then fails with an error instead of
// corrupting memory).  compute-sanitizer is closed on the pool this was developed on, so the
// gather tests are also run against a driver built with this switch (tests/test_header_api.py).
#ifdef SCLX_GX_CHECKS
#define SCLX_GX_ASSERT(cond)                                                                     \
    do {                                                                                         \
        if (!(cond)) {                                                                           \
            printf("gather_exchange.cuh:%d: check failed: %s\n", __LINE__, #cond);               \
            __trap();                                                                            \
        }                                                                                        \
    } while (0)
#else
#define SCLX_GX_ASSERT(cond) ((void)0)
#endif

namespace sclx::b200 {

constexpr int kGxMaxParts     = 16;   // devices on either side
constexpr int kGxMaxBuckets   = 512;  // owners x source regions
constexpr int kGxThreads      = 512;
constexpr int kGxPerThread    = 8;
constexpr int kGxTile         = kGxThreads * kGxPerThread;  // 4096 indices per tile
constexpr unsigned kGxRunAlign = 8;   // queue runs start on 32-byte sectors
constexpr size_t kGxPartitionSmem = size_t(kGxTile) * (sizeof(unsigned long long) + sizeof(unsigned));  // K1, dynamic

// e / per and e % per with one 64x64 high multiply (per is the same for every partition but
// the last: ceil(last/ndev) * leading elements)
struct gx_div_t {
    unsigned long long per, magic;
};
__host__ inline gx_div_t gx_make_div(unsigned long long per) {
    return gx_div_t{per, per ? ~0ull / per : 0ull};
}
__host__ __device__ __forceinline__ void
gx_owner(unsigned long long e, const gx_div_t& d, unsigned& q, unsigned& off) {
#ifdef __CUDA_ARCH__
    unsigned long long qq = __umul64hi(e, d.magic);
#else
    unsigned long long qq = (unsigned long long)(((unsigned __int128)e * d.magic) >> 64);
#endif
    unsigned long long r = e - qq * d.per;
    if (r >= d.per) {
        ++qq;
        r -= d.per;
    }
    q   = unsigned(qq);
    off = unsigned(r);
}

// source element -> bucket: owner q (uniform blocks of `per` elements), region s inside the
// owner's partition (2^region_shift elements each); bucket = q * regions + s
struct gx_geometry_t {
    gx_div_t div;
    unsigned magic32;  // floor(2^32 / per) when per < 2^32 (kGxOwner32)
    unsigned region_shift;
    unsigned regions;  // per owner
    unsigned buckets;  // nowners * regions
};

template<class U>  // U: unsigned carrier of the element width
struct gx_requester_t {
    U* out;             // this partition of `out`
    const U* src_base;  // element 0 of the whole source (direct path when a queue is full)
    size_t n;           // indices in this partition
    int nowners;
    gx_geometry_t geo;
    unsigned* queue[kGxMaxParts];  // (this requester -> owner q): [regions][cap] offsets, in q's memory
    U* resp[kGxMaxParts];          // [regions][cap] responses from owner q, local
    unsigned cap;                  // capacity of every bucket's queue (multiple of kGxRunAlign)
    unsigned* counters;            // [buckets] entries reserved so far, local
    unsigned* counters_pub[kGxMaxParts];  // [regions] copy of owner q's share of `counters`, in q's memory
    unsigned* ctas_done;           // local, zero-initialised, self-resetting
    unsigned short* pos;           // [tiles * kGxTile] in-tile positions of queued entries
    unsigned* tile_base;           // [tiles][buckets] first queue slot of the tile's run
    unsigned short* tile_cnt;      // [tiles][buckets] length of that run
};

template<class U>
struct gx_owner_t {
    const U* src_local;  // first element of this partition of the source
    size_t part_elems;   // elements in it
    int nreq;
    unsigned regions, region_shift;
    const unsigned* queue[kGxMaxParts];    // (r -> this owner): [regions][cap[r]] offsets, local
    const unsigned* counter[kGxMaxParts];  // [regions] fill levels of requester r's queues, published here by K1
    unsigned cap[kGxMaxParts];
    U* resp[kGxMaxParts];  // [regions][cap[r]] response buffer, in requester r's memory
};

// exclusive prefix sum of cnt[0..n) into start[0..n], by warp 0 of the CTA (n <= kGxMaxBuckets)
__device__ __forceinline__ void gx_scan_counts(const unsigned short* cnt, unsigned* start, int n) {
    if (threadIdx.x >= 32) return;
    const int lane = threadIdx.x;
    const int per  = (n + 31) / 32;
    const int lo = lane * per, hi = lo + per < n ? lo + per : n;
    unsigned sum = 0;
    for (int b = lo; b < hi; ++b) sum += cnt[b];
    unsigned incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
    }
    unsigned run = incl - sum;
    for (int b = lo; b < hi; ++b) {
        start[b] = run;
        run += cnt[b];
    }
    if (lane == 31) start[n] = incl;
}

// ---- K1 -----------------------------------------------------------------------------------
// Per-thread state is packed (32-bit offset + bucket<<16|rank) so three 512-thread CTAs fit an
// SM (profiles/r01_ncu_gather_exchange_details.txt: keeping the 64-bit indices live cost a CTA
// per SM and made the kernel issue-bound).
// how K1 finds the owner of a source element: compiled in, so the common cases carry no
// 64-bit division code (47 instructions per index in the generic form)
enum gx_owner_mode : int {
    kGxOneOwner   = 0,  // a single partition: the element index IS the offset
    kGxOwner32    = 1,  // several owners, fewer than 2^32 source elements: 32-bit magic division
    kGxOwner64    = 2   // general
};

// e / per for e < 2^32: one 32x32 high multiply and a correction step
__device__ __forceinline__ void gx_owner32(unsigned e, unsigned per, unsigned magic, unsigned& q, unsigned& off) {
    unsigned qq = __umulhi(e, magic);
    unsigned r  = e - qq * per;
    if (r >= per) {
        ++qq;
        r -= per;
    }
    q   = qq;
    off = r;
}

template<class U, class I, int OWNER_MODE>
__global__ void __launch_bounds__(kGxThreads, 3)
    gx_partition_kernel(gx_requester_t<U> rq, const I* __restrict__ index) {
    // dynamic shared memory (kGxPartitionSmem bytes):
    //   s_rec  one 64-bit record per queued index, in slot order: offset | bucket << 32 |
    //          position << 48 (one store when the tile is bucketed, one load when it is written)
    //   s_off  the offsets in index order between the two passes over the tile (keeping them in
    //          registers across three barriers spilled to local memory)
    extern __shared__ __align__(16) unsigned char s_dyn[];
    unsigned long long* s_rec = reinterpret_cast<unsigned long long*>(s_dyn);
    unsigned* s_off           = reinterpret_cast<unsigned*>(s_dyn + sizeof(unsigned long long) * kGxTile);
    __shared__ unsigned s_cnt[kGxMaxBuckets];
    __shared__ unsigned s_span[kGxMaxBuckets + 1];  // first slot | entries taken << 16, per bucket
    __shared__ unsigned s_start[kGxMaxBuckets + 1];
    __shared__ unsigned* s_run[kGxMaxBuckets];  // where the tile's run of each bucket starts in its queue
    __shared__ unsigned short s_take[kGxMaxBuckets];
    const unsigned tid     = threadIdx.x;
    const unsigned nb      = rq.geo.buckets;
    const unsigned regions = rq.geo.regions;
    const size_t tiles     = (rq.n + kGxTile - 1) / kGxTile;

    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        for (unsigned b = tid; b < nb; b += kGxThreads) s_cnt[b] = 0;
        __syncthreads();
        const size_t tbase   = t * kGxTile;
        const unsigned avail = unsigned(rq.n - tbase < size_t(kGxTile) ? rq.n - tbase : size_t(kGxTile));
        const I* __restrict__ tile_index = index + tbase;
        unsigned br[kGxPerThread];  // bucket << 16 | rank in (tile, bucket)
        {
            unsigned long long e[kGxPerThread];
#pragma unroll
            for (int k = 0; k < kGxPerThread; ++k) {
                const unsigned j = unsigned(k) * kGxThreads + tid;
                e[k]             = j < avail ? (unsigned long long)__ldg(tile_index + j) : ~0ull;
            }
#pragma unroll
            for (int k = 0; k < kGxPerThread; ++k) {
                const bool valid = unsigned(k) * kGxThreads + tid < avail;
                unsigned b       = 0xffffu, off = 0;
                if (valid) {
                    unsigned q = 0;
                    if constexpr (OWNER_MODE == kGxOneOwner) off = unsigned(e[k]);
                    else if constexpr (OWNER_MODE == kGxOwner32)
                        gx_owner32(unsigned(e[k]), unsigned(rq.geo.div.per), rq.geo.magic32, q, off);
                    else gx_owner(e[k], rq.geo.div, q, off);
                    if (OWNER_MODE != kGxOneOwner && q >= unsigned(rq.nowners))
                        q = unsigned(rq.nowners) - 1;  // out of range: UB upstream too
                    unsigned s = off >> rq.geo.region_shift;
                    if (s >= regions) s = regions - 1;
                    b = q * regions + s;
                }
                s_off[unsigned(k) * kGxThreads + tid] = off;
                // rank inside (tile, bucket): one shared-memory atomic per index.  With hundreds
                // of buckets the lanes of a warp rarely collide; aggregating equal buckets with
                // __match_any_sync first cost 3x the whole kernel (profiles/r02_ncu_gather_binned_v1.txt)
                br[k] = valid ? ((b << 16) | atomicAdd(&s_cnt[b], 1u)) : 0xffffffffu;
            }
        }
        __syncthreads();
        for (unsigned b = tid; b < nb; b += kGxThreads) {
            // runs are reserved in multiples of kGxRunAlign entries: every run starts on a
            // 32-byte sector of the queue and of the response buffer
            const unsigned c    = s_cnt[b];
            const unsigned g    = c ? atomicAdd(&rq.counters[b], (c + kGxRunAlign - 1u) & ~(kGxRunAlign - 1u)) : 0u;
            const unsigned room = g < rq.cap ? rq.cap - g : 0u;
            const unsigned take = c < room ? c : room;
            const unsigned q    = b / regions, sub = b - q * regions;
            s_run[b]            = rq.queue[q] + size_t(sub) * rq.cap + g;
            s_take[b]           = (unsigned short)take;
            rq.tile_base[t * nb + b] = g;
            rq.tile_cnt[t * nb + b]  = (unsigned short)take;
        }
        __syncthreads();
        gx_scan_counts(s_take, s_start, int(nb));
        __syncthreads();
        for (unsigned b = tid; b < nb; b += kGxThreads) s_span[b] = s_start[b] | (unsigned(s_take[b]) << 16);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kGxPerThread; ++k) {
            const unsigned j = unsigned(k) * kGxThreads + tid;
            if (j < avail) {
                const unsigned b = br[k] >> 16, rank = br[k] & 0xffffu;
                const unsigned span = s_span[b];
                if (rank < (span >> 16)) {
                    SCLX_GX_ASSERT((span & 0xffffu) + rank < unsigned(kGxTile));
                    s_rec[(span & 0xffffu) + rank]
                        = (unsigned long long)s_off[j] | ((unsigned long long)b << 32) | ((unsigned long long)j << 48);
                } else {
                    // queue full: direct (peer) load
                    rq.out[tbase + j] = rq.src_base[(unsigned long long)__ldg(tile_index + j)];
                }
            }
        }
        __syncthreads();
        const unsigned total = s_start[nb];
        for (unsigned i = tid; i < total; i += kGxThreads) {
            // slots of one bucket are consecutive: a warp writes whole runs, coalesced
            const unsigned long long rec = s_rec[i];
            const unsigned b    = unsigned(rec >> 32) & 0xffffu;
            const unsigned span = s_span[b];
            SCLX_GX_ASSERT(b < nb && i >= (span & 0xffffu) && i - (span & 0xffffu) < (span >> 16));
            SCLX_GX_ASSERT(rq.tile_base[t * nb + b] + (i - (span & 0xffffu)) < rq.cap);
            s_run[b][i - (span & 0xffffu)] = unsigned(rec);
            rq.pos[tbase + i]              = (unsigned short)(rec >> 48);
        }
        // the padding slots of every run are served too: make them harmless (one thread per
        // bucket; inside the loop above the branch made every warp walk the padding path)
        for (unsigned b = tid; b < nb; b += kGxThreads) {
            const unsigned take = s_take[b];
            SCLX_GX_ASSERT(take == 0 || rq.tile_base[t * nb + b] + ((take + kGxRunAlign - 1u) & ~(kGxRunAlign - 1u)) <= rq.cap);
            for (unsigned p = take; p < ((take + kGxRunAlign - 1u) & ~(kGxRunAlign - 1u)); ++p) s_run[b][p] = 0u;
        }
        __syncthreads();
    }
    // The CTA that finishes last publishes the fill levels to the owners' memory: K2 then reads
    // a LOCAL word per CTA.  (Reading the requester's counter across NVLink from each of K2's
    // ~10^5 CTAs made the serve phase 2.8x slower per entry on two GPUs than on one.)
    __shared__ bool s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(rq.ctas_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (unsigned b = tid; b < nb; b += kGxThreads) {
        const unsigned q = b / regions, sub = b - q * regions;
        rq.counters_pub[q][sub] = __ldcg(rq.counters + b);
    }
    if (tid == 0) *rq.ctas_done = 0;
}

// ---- K2 -----------------------------------------------------------------------------------
template<class U>
__device__ __forceinline__ U gx_load_random(const U* p) {
    if constexpr (sizeof(U) == 4) {
        unsigned v;
        asm volatile("ld.global.nc.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
        return U(v);
    } else if constexpr (sizeof(U) == 8) {
        unsigned long long v;
        asm volatile("ld.global.nc.L2::64B.b64 %0, [%1];" : "=l"(v) : "l"(p));
        return U(v);
    } else {
        return *p;
    }
}

// sequential pull of [begin, end) into L2 (one instruction for the whole range)
__device__ __forceinline__ void gx_prefetch_l2(const void* begin, const void* end) {
    unsigned long long a0 = reinterpret_cast<unsigned long long>(begin) & ~15ull;
    unsigned long long a1 = (reinterpret_cast<unsigned long long>(end) + 15ull) & ~15ull;
    if (a1 <= a0) return;
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(unsigned(a1 - a0)) : "memory");
}

constexpr int kGxServeThreads = 256, kGxServeUnroll = 8;
constexpr int kGxServeChunk   = kGxServeThreads * kGxServeUnroll;

// grid = (ceil(max cap / chunk), nreq, regions): CTA (x, r, s) serves entries
// [x*chunk, (x+1)*chunk) of requester r's queue for region s of this owner.  CTAs are
// dispatched x-fastest, region-slowest, so the chip works on one region (plus the tail of the
// previous one) at a time; every CTA first pulls its slice of region s+1 into L2.  CTAs beyond
// a queue's fill level exit once their prefetch is issued.  One chunk per CTA and no loop:
// staggered CTA starts de-synchronise the queue-read / gather / store phases
// (experiments/serve_tune.cu, profiles/r01_serve_tune.txt).
template<class U>
__global__ void __launch_bounds__(kGxServeThreads) gx_serve_kernel(gx_owner_t<U> ow, int prefetch) {
    const unsigned s = blockIdx.z;
    const int r      = blockIdx.y;
    if (prefetch && threadIdx.x == 0) {
        // region 0 has no predecessor: its own CTAs pull it (late for some, still a stream)
        const unsigned target     = (s + 1 < ow.regions && prefetch == 1) ? s + 1 : s;
        const bool wanted         = target != s || s == 0;
        const size_t region_elems = size_t(1) << ow.region_shift;
        const size_t first        = size_t(target) * region_elems;
        if (wanted && first < ow.part_elems) {
            const size_t last   = first + region_elems < ow.part_elems ? first + region_elems : ow.part_elems;
            const size_t slices = size_t(gridDim.x) * gridDim.y;
            const size_t slice  = size_t(blockIdx.y) * gridDim.x + blockIdx.x;
            size_t step         = ((last - first + slices - 1) / slices + 31) & ~size_t(31);
            const size_t lo     = first + slice * step;
            if (lo < last) gx_prefetch_l2(ow.src_local + lo, ow.src_local + (lo + step < last ? lo + step : last));
        }
    }
    unsigned cnt = __ldcg(ow.counter[r] + s);
    if (cnt > ow.cap[r]) cnt = ow.cap[r];
    const unsigned base = blockIdx.x * unsigned(kGxServeChunk);
    if (base >= cnt) return;
    const unsigned* __restrict__ queue = ow.queue[r] + size_t(s) * ow.cap[r] + base;
    U* __restrict__ resp               = ow.resp[r] + size_t(s) * ow.cap[r] + base;
    const unsigned left                = cnt - base;
    unsigned o[kGxServeUnroll];
    U v[kGxServeUnroll];
#pragma unroll
    for (int u = 0; u < kGxServeUnroll; ++u) {
        const unsigned i = unsigned(u) * kGxServeThreads + threadIdx.x;
        o[u]             = i < left ? __ldcs(queue + i) : 0u;
    }
#pragma unroll
    for (int u = 0; u < kGxServeUnroll; ++u) {
        SCLX_GX_ASSERT(size_t(o[u]) < ow.part_elems);
        SCLX_GX_ASSERT(unsigned(u) * kGxServeThreads + threadIdx.x >= left
                       || (size_t(o[u]) >> ow.region_shift) == s || (s + 1 == ow.regions && (size_t(o[u]) >> ow.region_shift) >= s)
                       || o[u] == 0u);
        v[u] = gx_load_random(ow.src_local + o[u]);
    }
#pragma unroll
    for (int u = 0; u < kGxServeUnroll; ++u) {
        const unsigned i = unsigned(u) * kGxServeThreads + threadIdx.x;
        if (i < left) __stcs(resp + i, v[u]);
    }
}

// ---- K3 -----------------------------------------------------------------------------------
// One tile per CTA trip.  The tile's runs are laid out back to back in slot order (the order K1
// wrote `pos` in); the warps first expand the run table into a slot -> bucket map in shared
// memory, then every thread moves 8 slots with 8 independent loads in flight: response value
// (coalesced inside a run) + position.  The values are un-permuted in SHARED memory and the
// tile's window of `out` is written with full-width coalesced stores (one scattered 4-byte
// store per index is one L2 request per index: profiles/r02_gather_v2_launches.csv).  Indices
// K1 served directly (queue overflow) are already in `out` and are left alone.
template<class U>
__global__ void __launch_bounds__(kGxThreads, 3) gx_place_kernel(gx_requester_t<U> rq) {
    __shared__ U s_val[kGxTile];
    __shared__ unsigned s_start[kGxMaxBuckets + 1];
    __shared__ const U* s_run[kGxMaxBuckets];
    __shared__ unsigned short s_take[kGxMaxBuckets], s_bkt[kGxTile];
    __shared__ unsigned s_written[kGxTile / 32];  // bit per window slot: value came through a queue
    const unsigned tid     = threadIdx.x;
    const unsigned nb      = rq.geo.buckets;
    const unsigned regions = rq.geo.regions;
    const size_t tiles     = (rq.n + kGxTile - 1) / kGxTile;
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        for (unsigned b = tid; b < nb; b += kGxThreads) {
            const unsigned q = b / regions, sub = b - q * regions;
            s_run[b]  = rq.resp[q] + size_t(sub) * rq.cap + __ldcs(rq.tile_base + t * nb + b);
            s_take[b] = __ldcs(rq.tile_cnt + t * nb + b);
        }
        if (tid < kGxTile / 32) s_written[tid] = 0;
        __syncthreads();
        gx_scan_counts(s_take, s_start, int(nb));
        __syncthreads();
        // slot -> bucket map: one thread per bucket stamps its run (a warp per bucket with a
        // lane loop cost half of the kernel's instructions: profiles/r02_ncu_gather_binned_v4.txt)
        for (unsigned b = tid; b < nb; b += kGxThreads) {
            const unsigned first = s_start[b], take = s_take[b];
            for (unsigned j = 0; j < take; ++j) s_bkt[first + j] = (unsigned short)b;
        }
        __syncthreads();
        const size_t tbase   = t * kGxTile;
        const unsigned avail = unsigned(rq.n - tbase < size_t(kGxTile) ? rq.n - tbase : size_t(kGxTile));
        const unsigned total = s_start[nb];
        U v[kGxPerThread];
        unsigned short p[kGxPerThread];
#pragma unroll
        for (int k = 0; k < kGxPerThread; ++k) {
            const unsigned i = unsigned(k) * kGxThreads + tid;
            if (i < total) {
                const unsigned b = s_bkt[i];
                SCLX_GX_ASSERT(b < nb && i >= s_start[b] && i - s_start[b] < s_take[b]);
                v[k]             = __ldcs(s_run[b] + (i - s_start[b]));
                p[k]             = __ldcs(rq.pos + tbase + i);
            }
        }
#pragma unroll
        for (int k = 0; k < kGxPerThread; ++k) {
            const unsigned i = unsigned(k) * kGxThreads + tid;
            if (i < total) {
                SCLX_GX_ASSERT(p[k] < avail);
                s_val[p[k]] = v[k];
                if (total != avail) atomicOr(&s_written[p[k] >> 5], 1u << (p[k] & 31));
            }
        }
        __syncthreads();
        U* __restrict__ window = rq.out + tbase;
        if (total == avail) {
            // the common case: every index of the tile went through a queue
            for (unsigned j = tid; j < avail; j += kGxThreads) window[j] = s_val[j];
        } else {
            for (unsigned j = tid; j < avail; j += kGxThreads)
                if (s_written[j >> 5] >> (j & 31) & 1u) window[j] = s_val[j];
        }
        __syncthreads();
    }
}

// direct form (one partition of `out`): every thread keeps UNROLL independent loads in flight;
// `src` is element 0 of the whole source, so cross-partition indices are peer loads
template<class U, class I, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS)
    gather_direct_kernel(const I* __restrict__ index, size_t n, const U* __restrict__ src, U* __restrict__ out) {
    constexpr size_t CHUNK = size_t(THREADS) * UNROLL;  // contiguous per CTA and trip
    for (size_t base = size_t(blockIdx.x) * CHUNK; base < n; base += size_t(gridDim.x) * CHUNK) {
        const size_t left = n - base;
        unsigned long long j[UNROLL];
        U v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const size_t i = size_t(u) * THREADS + threadIdx.x;
            j[u]           = i < left ? (unsigned long long)__ldg(index + base + i) : 0ull;
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const size_t i = size_t(u) * THREADS + threadIdx.x;
            if (i < left) v[u] = gx_load_random(src + j[u]);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const size_t i = size_t(u) * THREADS + threadIdx.x;
            if (i < left) out[base + i] = v[u];
        }
    }
}

}  // namespace sclx::b200
