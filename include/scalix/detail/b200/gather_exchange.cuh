// scalix-b200: exchange form of the cross-partition gather  out[i] = src[index[i]].
//
// Why it exists (experiments/p2p_probe.cu on 2 x B200, profiles/r01_p2p_probe.txt): NVLink 5
// P2P serves about 6.5 G REQUESTS/s per GPU whatever the request carries — a random 4-byte
// peer load, a 16-byte one, a posted 4-byte store and a fully coalesced 128-byte line all cost
// one request.  A gather that issues one peer load per index (what a user lambda over the
// range launcher does, and all an opaque lambda can do) is therefore capped at ~26 GB/s of
// payload, 3 % of the link, while local HBM serves 40 G random 4-byte reads/s.  The exchange
// form moves the same data in 128-byte requests only:
//
//   K1  partition (requester r, one CTA per 4096-index tile): classify every index by owning
//       partition q, bucket the tile in shared memory, reserve a run in the (r -> q) request
//       queue with ONE atomic per bucket and write the 32-bit in-partition offsets to the queue
//       IN OWNER q's MEMORY with coalesced posted stores; keep the 16-bit in-tile positions
//       locally, in the same order.
//   K2  serve (owner q): stream its queues, gather from its LOCAL partition (HBM random reads)
//       and write the values, in queue order, to the response buffer IN REQUESTER r's MEMORY
//       (coalesced posted stores).
//   K3  place (requester r): stream responses + positions, scatter inside the tile's 16 KiB
//       window of `out`.
//
// The three kernels are three launch groups (ordered across devices by stream events, no host
// round trip).  Queues have bounded capacity (1.25 x the uniform share + one tile); an index
// that finds its queue full is served by a direct peer load in K1, so skewed index sets stay
// correct and merely lose the speed-up.  Pure data movement: results are bit-identical to the
// direct gather whatever the order the queues were filled in.
#pragma once

#include "primitives.cuh"

#include <cstdint>

namespace sclx::b200 {

constexpr int kGxMaxParts  = 16;
constexpr int kGxThreads   = 512;
constexpr int kGxPerThread = 8;
constexpr int kGxTile      = kGxThreads * kGxPerThread;  // 4096 indices per tile

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

template<class U>  // U: unsigned carrier of the element width
struct gx_requester_t {
    U* out;             // this partition of `out`
    const U* src_base;  // element 0 of the whole source (direct path when a queue is full)
    size_t n;           // indices in this partition
    int nowners;
    gx_div_t div;
    unsigned* queue[kGxMaxParts];  // (this requester -> q) offsets, in owner q's memory
    unsigned cap[kGxMaxParts];     // capacity of that queue / response buffer (multiple of 32)
    U* resp[kGxMaxParts];          // responses from owner q, local
    unsigned* counters;            // [nowners] entries reserved so far, local
    unsigned short* pos;           // [tiles * kGxTile] in-tile positions of queued entries
    unsigned* tile_base;           // [tiles][nowners] first queue slot of the tile's run
    unsigned short* tile_cnt;      // [tiles][nowners] length of that run
};

template<class U>
struct gx_owner_t {
    const U* src_local;  // first element of this partition of the source
    int nreq;
    const unsigned* queue[kGxMaxParts];    // (r -> this owner) offsets, local
    const unsigned* counter[kGxMaxParts];  // &counters_r[this owner], in requester r's memory
    unsigned cap[kGxMaxParts];
    U* resp[kGxMaxParts];  // response buffer for this owner, in requester r's memory
};

// ---- K1 -----------------------------------------------------------------------------------
// Per-thread state is packed (32-bit offset + owner<<16|rank) so three 512-thread CTAs fit an
// SM; the first version kept the 64-bit indices live, needed 58 registers, ran two CTAs per SM
// and was issue-bound at 1.0 ms for 2^27 indices (profiles/r01_ncu_gather_exchange_details.txt).
template<class U, class I>
__global__ void __launch_bounds__(kGxThreads, 3)
    gx_partition_kernel(gx_requester_t<U> rq, const I* __restrict__ index) {
    __shared__ unsigned s_off[kGxTile];
    __shared__ unsigned short s_pos[kGxTile];
    __shared__ unsigned s_cnt[kGxMaxParts], s_start[kGxMaxParts + 1], s_gbase[kGxMaxParts],
        s_take[kGxMaxParts];
    const unsigned tid   = threadIdx.x;
    const unsigned lane  = tid & 31;
    const unsigned below = (1u << lane) - 1u;
    const int nown       = rq.nowners;
    const size_t tiles   = (rq.n + kGxTile - 1) / kGxTile;

    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        if (tid < unsigned(nown)) s_cnt[tid] = 0;
        __syncthreads();
        const size_t tbase   = t * kGxTile;
        const unsigned avail = unsigned(rq.n - tbase < size_t(kGxTile) ? rq.n - tbase : size_t(kGxTile));
        const I* __restrict__ tile_index = index + tbase;
        unsigned off[kGxPerThread], qr[kGxPerThread];  // qr = owner << 16 | rank in (tile, owner)
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
                unsigned q       = 0xffu;
                off[k]           = 0;
                if (valid) {
                    gx_owner(e[k], rq.div, q, off[k]);
                    if (q >= unsigned(nown)) q = unsigned(nown) - 1;  // out of range: UB upstream too
                }
                // one shared-memory atomic per (warp, owner)
                const unsigned peers = __match_any_sync(0xffffffffu, q);
                const int leader     = __ffs(peers) - 1;
                unsigned first       = 0;
                if (int(lane) == leader && valid) first = atomicAdd(&s_cnt[q], unsigned(__popc(peers)));
                first = __shfl_sync(0xffffffffu, first, leader);
                qr[k] = (q << 16) | (first + unsigned(__popc(peers & below)));
            }
        }
        __syncthreads();
        if (tid < unsigned(nown)) {
            // runs are reserved in multiples of 32 entries, so every run starts on a 128-byte
            // line of the queue and a warp's store of 32 offsets is one full-line NVLink write
            const unsigned c    = s_cnt[tid];
            const unsigned g    = c ? atomicAdd(&rq.counters[tid], (c + 31u) & ~31u) : 0u;
            const unsigned room = g < rq.cap[tid] ? rq.cap[tid] - g : 0u;
            const unsigned take = c < room ? c : room;
            s_gbase[tid]        = g;
            s_take[tid]         = take;
            rq.tile_base[t * nown + tid] = g;
            rq.tile_cnt[t * nown + tid]  = (unsigned short)take;
        }
        __syncthreads();
        if (tid == 0) {
            unsigned acc = 0;
            for (int o = 0; o < nown; ++o) {
                s_start[o] = acc;
                acc += s_take[o];
            }
            s_start[nown] = acc;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kGxPerThread; ++k) {
            const unsigned j = unsigned(k) * kGxThreads + tid;
            if (j < avail) {
                const unsigned q = qr[k] >> 16, rank = qr[k] & 0xffffu;
                if (rank < s_take[q]) {
                    const unsigned slot = s_start[q] + rank;
                    s_off[slot]         = off[k];
                    s_pos[slot]         = (unsigned short)j;
                } else {
                    // queue full: direct (peer) load
                    rq.out[tbase + j] = rq.src_base[(unsigned long long)__ldg(tile_index + j)];
                }
            }
        }
        __syncthreads();
        const unsigned total = s_start[nown];
        for (int o = 0; o < nown; ++o) {
            const unsigned take = s_take[o], padded = (take + 31u) & ~31u;  // padding reads offset 0
            unsigned* __restrict__ queue = rq.queue[o] + s_gbase[o];
            const unsigned first         = s_start[o];
            for (unsigned j = tid; j < padded; j += kGxThreads) queue[j] = j < take ? s_off[first + j] : 0u;
        }
        for (unsigned i = tid; i < total; i += kGxThreads) rq.pos[tbase + i] = s_pos[i];
        __syncthreads();
    }
}

// ---- K2 -----------------------------------------------------------------------------------
// One random 4/8-byte load per entry: bound by the HBM's random-access rate, not by bytes.
// experiments/serve_tune.cu (profiles/r01_serve_tune.txt): a persistent grid in lock step
// serves 36-37 G loads/s, many small CTAs (one 1024-entry chunk each, no loop) 47-48 G/s —
// the staggered CTA starts de-synchronise the queue-read / random-read / store phases.  The
// L2::64B hint halves the DRAM bytes fetched per load (126 -> 63).
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

constexpr int kGxServeThreads = 256, kGxServeUnroll = 4;
constexpr int kGxServeChunk   = kGxServeThreads * kGxServeUnroll;

// grid = (ceil(max cap / chunk), nreq): CTA (x, r) serves entries [x*chunk, (x+1)*chunk) of
// requester r's queue; CTAs beyond the queue's fill level exit at once
template<class U>
__global__ void __launch_bounds__(kGxServeThreads) gx_serve_kernel(gx_owner_t<U> ow) {
    const int r  = blockIdx.y;
    unsigned cnt = __ldcg(ow.counter[r]);
    if (cnt > ow.cap[r]) cnt = ow.cap[r];
    const unsigned base = blockIdx.x * unsigned(kGxServeChunk);
    if (base >= cnt) return;
    const unsigned* __restrict__ queue = ow.queue[r] + base;
    U* __restrict__ resp               = ow.resp[r] + base;
    const unsigned left                = cnt - base;
    unsigned o[kGxServeUnroll];
    U v[kGxServeUnroll];
#pragma unroll
    for (int u = 0; u < kGxServeUnroll; ++u) {
        const unsigned i = unsigned(u) * kGxServeThreads + threadIdx.x;
        o[u]             = i < left ? __ldcs(queue + i) : 0u;
    }
#pragma unroll
    for (int u = 0; u < kGxServeUnroll; ++u) v[u] = gx_load_random(ow.src_local + o[u]);
#pragma unroll
    for (int u = 0; u < kGxServeUnroll; ++u) {
        const unsigned i = unsigned(u) * kGxServeThreads + threadIdx.x;
        if (i < left) resp[i] = v[u];
    }
}

// ---- K3 -----------------------------------------------------------------------------------
template<class U>
__global__ void __launch_bounds__(kGxThreads) gx_place_kernel(gx_requester_t<U> rq) {
    __shared__ unsigned s_start[kGxMaxParts + 1], s_gbase[kGxMaxParts];
    const int tid      = threadIdx.x;
    const int nown     = rq.nowners;
    const size_t tiles = (rq.n + kGxTile - 1) / kGxTile;
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        if (tid == 0) {
            unsigned acc = 0;
            for (int o = 0; o < nown; ++o) {
                s_start[o] = acc;
                s_gbase[o] = rq.tile_base[t * nown + o];
                acc += rq.tile_cnt[t * nown + o];
            }
            s_start[nown] = acc;
        }
        __syncthreads();
        const size_t tbase   = t * kGxTile;
        const unsigned total = s_start[nown];
        U v[kGxPerThread];
        unsigned short p[kGxPerThread];
#pragma unroll
        for (int k = 0; k < kGxPerThread; ++k) {
            const unsigned i = unsigned(k) * kGxThreads + tid;
            if (i < total) {
                int o = 0;
                while (i >= s_start[o + 1]) ++o;
                v[k] = __ldcs(rq.resp[o] + s_gbase[o] + (i - s_start[o]));
                p[k] = __ldcs(rq.pos + tbase + i);
            }
        }
#pragma unroll
        for (int k = 0; k < kGxPerThread; ++k) {
            const unsigned i = unsigned(k) * kGxThreads + tid;
            if (i < total) rq.out[tbase + p[k]] = v[k];
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
