// scalix-b200: reductions for sm_100a.
//
//  * reduce_kernel            whole-slice fold to one value (one pass, 128-bit loads).
//                             Replaces thrust::reduce at
//                             /root/reference scalix/algorithm/reduce.cuh:73-79 and is
//                             pass A (per-partition total) of the multi-device scan.
//  * reduce_last_dim_kernel   out[s] = fold_j in[s + stride*j] for ALL rows s in one
//                             pass that reads every byte once, coalesced.  Replaces the
//                             `stride` serial strided thrust::reduce calls at
//                             /root/reference scalix/algorithm/reduce_last_dim.cuh:189-206.
//  * fold_rows_kernel         deterministic fixed-order fold of per-partition rows
//                             (replaces the mutex-ordered host fold, ibid. :208-211).
//
// All folds run in a fixed order that depends only on (shape, grid), so float
// results are bit-stable run to run.  Operators must be associative and
// commutative (as thrust::reduce requires of the reference's callers).
#pragma once

#include "primitives.cuh"

namespace sclx::b200 {

// block-wide fold; result valid in thread 0.  `smem` holds THREADS/32 values.
template<class T, class Op, int THREADS>
__device__ __forceinline__ T block_reduce(T x, Op& op, T* smem) {
    constexpr int WARPS = THREADS / 32;
    const int lane      = threadIdx.x & 31;
    const int warp      = threadIdx.x >> 5;
    x                   = warp_reduce(x, op);
    if (lane == 0) smem[warp] = x;
    __syncthreads();
    if (warp == 0) {
        T y = smem[lane < WARPS ? lane : 0];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            T z = shfl_down(y, d);
            if (lane + d < WARPS) y = op(y, z);
        }
        x = y;
    }
    return x;
}

struct reduce_workspace_t {
    void* partials;          // gridDim.x values (reduce) or gridDim.x*rows (last_dim)
    unsigned* ticket;        // zero-initialised, self-resetting
};

// Cross-partition step of reduce_last_dim fused into the kernel (replaces the mutex-ordered
// host fold of /root/reference scalix/algorithm/reduce_last_dim.cuh:208-211 AND a second launch
// group): every partition's kernel stores its row into peer-visible memory `rows`
// ([nparts][row length], plain NVLink stores), then bumps `arrivals` with a system-scope
// atomic; the partition that arrives LAST folds all rows in fixed partition order into
// `final_out`.  Nobody waits for anybody, so it is deadlock-free whatever the devices'
// relative timing (or when N logical devices share one GPU).  nparts <= 1: no exchange.
struct row_exchange_t {
    void* rows         = nullptr;  // [nparts][rows], peer-visible
    unsigned* arrivals = nullptr;  // zero-initialised, wraps back to zero (atomicInc)
    void* final_out    = nullptr;  // rows values, peer-visible
    int nparts         = 0;
    int ordinal        = 0;
    // optional: pinned, mapped host word that receives `done_ticket` once `final_out` is
    // written — the caller polls it instead of synchronising every device's stream
    unsigned* done_flag  = nullptr;
    unsigned done_ticket = 0;
};

// ---------------------------------------------------------------------------
// whole-slice reduce
// ---------------------------------------------------------------------------
template<class T, class Op, int THREADS, int VECS, bool ALIGNED>
__global__ void __launch_bounds__(THREADS) reduce_kernel(
    const T* __restrict__ in,
    size_t n,
    T identity,
    T* __restrict__ out,  // one value; receives op(identity-seeded fold)
    reduce_workspace_t ws,
    Op op
) {
    constexpr int V = vec_traits<T>::width;
    __shared__ T s_red[THREADS / 32];
    __shared__ bool s_last;

    T acc[VECS];
#pragma unroll
    for (int k = 0; k < VECS; ++k) acc[k] = identity;

    const size_t nvec        = ALIGNED ? n / V : 0;
    const size_t step        = size_t(gridDim.x) * THREADS;
    const size_t first       = size_t(blockIdx.x) * THREADS + threadIdx.x;
    if constexpr (ALIGNED && V > 1) {
        size_t i = first;
        // VECS independent 128-bit loads in flight per thread
        for (; i + (VECS - 1) * step < nvec; i += VECS * step) {
            vec_t<T, V> v[VECS];
#pragma unroll
            for (int k = 0; k < VECS; ++k)
                v[k] = load_vec_stream<T, V>(in + (i + k * step) * V);
#pragma unroll
            for (int k = 0; k < VECS; ++k) {
#pragma unroll
                for (int e = 0; e < V; ++e) acc[k] = op(acc[k], v[k].v[e]);
            }
        }
        for (; i < nvec; i += step) {
            vec_t<T, V> v = load_vec_stream<T, V>(in + i * V);
#pragma unroll
            for (int e = 0; e < V; ++e) acc[0] = op(acc[0], v.v[e]);
        }
    }
    // scalar tail (or the whole range when unaligned)
    for (size_t i = nvec * V + first; i < n; i += step) acc[0] = op(acc[0], in[i]);

#pragma unroll
    for (int k = 1; k < VECS; ++k) acc[0] = op(acc[0], acc[k]);

    T block = block_reduce<T, Op, THREADS>(acc[0], op, s_red);

    T* partials = static_cast<T*>(ws.partials);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = block;
        __threadfence();
        unsigned t = atomicAdd(ws.ticket, 1u);
        s_last     = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // last CTA: fixed-order fold of the per-CTA partials
    T a = identity;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += THREADS)
        a = op(a, load_cg(partials + i));
    __syncthreads();
    a = block_reduce<T, Op, THREADS>(a, op, s_red);
    if (threadIdx.x == 0) {
        *out       = a;
        *ws.ticket = 0;  // ready for the next launch on this stream
    }
}

// ---------------------------------------------------------------------------
// reduce along the last (slowest, partitioned) dimension
// ---------------------------------------------------------------------------
// Memory is [cols][rows] with rows contiguous (rows = product of the leading
// dims = the reference's `stride`).  Work unit = one vector of VW elements
// (VW = 16/sizeof(T) when rows % VW == 0 and the base is 16-byte aligned, else 1).
// slots = rows / VW vectors per column.
//   slots <= THREADS : `active` = floor(THREADS/slots)*slots threads, thread t keeps
//                      slot t % slots for its whole life and walks columns, so its
//                      VW accumulators live in registers; consecutive threads read
//                      consecutive vectors (fully coalesced).
//   slots  > THREADS : blockIdx.y selects a band of THREADS slots.
// UNROLL independent vector loads are in flight per thread; they fold into ACCS accumulator
// sets (ACCS < UNROLL frees registers for more resident CTAs: the loads carry the latency, the
// adds do not).
template<class T, class Op, int THREADS, int UNROLL, int VW, int ACCS = UNROLL>
__global__ void __launch_bounds__(THREADS) reduce_last_dim_kernel(
    const T* __restrict__ in,
    size_t rows,
    size_t cols,
    T identity,
    T* __restrict__ out,  // rows values (ignored when xchg.nparts > 1)
    reduce_workspace_t ws,
    Op op,
    row_exchange_t xchg = row_exchange_t{}
) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    T* s_acc = reinterpret_cast<T*>(s_raw);  // [groups][slots_in_band*VW]
    __shared__ bool s_last;

    const size_t slots = rows / VW;
    const int tid      = threadIdx.x;

    size_t band_first, band_slots;  // slots handled by this CTA
    int groups;                     // column groups processed side by side
    if (slots <= THREADS) {
        band_first = 0;
        band_slots = slots;
        groups     = int(THREADS / slots);
    } else {
        band_first = size_t(blockIdx.y) * THREADS;
        band_slots = (slots - band_first < THREADS) ? slots - band_first : THREADS;
        groups     = 1;
    }
    const int active = int(groups * band_slots);
    const int slot   = int(tid % band_slots);
    const int group  = int(tid / band_slots);

    static_assert(ACCS >= 1 && ACCS <= UNROLL && (ACCS & (ACCS - 1)) == 0, "ACCS: power of two <= UNROLL");
    T acc[ACCS][VW];
#pragma unroll
    for (int u = 0; u < ACCS; ++u)
#pragma unroll
        for (int e = 0; e < VW; ++e) acc[u][e] = identity;

    if (tid < active) {
        const size_t col_step = size_t(gridDim.x) * groups;
        size_t c              = size_t(blockIdx.x) * groups + group;
        const T* base         = in + (band_first + slot) * VW;
        for (; c + (UNROLL - 1) * col_step < cols; c += UNROLL * col_step) {
            if constexpr (VW > 1) {
                vec_t<T, VW> v[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    v[u] = load_vec_stream<T, VW>(base + (c + u * col_step) * rows);
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
#pragma unroll
                    for (int e = 0; e < VW; ++e) acc[u % ACCS][e] = op(acc[u % ACCS][e], v[u].v[e]);
            } else {
                T v[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    v[u] = __ldg(base + (c + u * col_step) * rows);
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) acc[u % ACCS][0] = op(acc[u % ACCS][0], v[u]);
            }
        }
        for (; c < cols; c += col_step) {
#pragma unroll
            for (int e = 0; e < VW; ++e) acc[0][e] = op(acc[0][e], base[c * rows + e]);
        }
        // pairwise fold of the unrolled accumulators
#pragma unroll
        for (int w = 1; w < ACCS; w <<= 1)
#pragma unroll
            for (int u = 0; u + w < ACCS; u += 2 * w)
#pragma unroll
                for (int e = 0; e < VW; ++e) acc[u][e] = op(acc[u][e], acc[u + w][e]);
#pragma unroll
        for (int e = 0; e < VW; ++e)
            s_acc[(size_t(group) * band_slots + slot) * VW + e] = acc[0][e];
    }
    __syncthreads();

    // fold the column groups (tree over groups, fixed order), then publish
    const size_t band_elems = band_slots * VW;
    T* partials             = static_cast<T*>(ws.partials);
    for (size_t r = tid; r < band_elems; r += THREADS) {
        T a = s_acc[r];
        for (int g = 1; g < groups; ++g) a = op(a, s_acc[size_t(g) * band_elems + r]);
        partials[size_t(blockIdx.x) * rows + band_first * VW + r] = a;
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        unsigned t = atomicAdd(ws.ticket, 1u);
        s_last     = (t == gridDim.x * gridDim.y - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // last CTA: out[r] = identity (+) partial[0][r] (+) partial[1][r] ... in a FIXED order.
    // All threads take part: with rows <= THREADS the CTA range is split into THREADS/rows
    // contiguous pieces folded side by side (8 independent chains each, so ~8 L2 loads are in
    // flight per thread), then the pieces are folded in order.  (One thread per row walking
    // every CTA's partial cost ~45 us of a 670 us kernel at 592 CTAs.)
    const unsigned nblk = gridDim.x;
    const bool exchange = xchg.nparts > 1;
    if (exchange) out = static_cast<T*>(xchg.rows) + size_t(xchg.ordinal) * rows;
    if (rows <= size_t(THREADS)) {
        const unsigned pieces = unsigned(THREADS / rows);
        const unsigned chunk  = (nblk + pieces - 1) / pieces;
        __syncthreads();
        if (size_t(tid) < size_t(pieces) * rows) {
            const size_t r       = size_t(tid) % rows;
            const unsigned piece = unsigned(size_t(tid) / rows);
            unsigned b           = piece * chunk;
            const unsigned end   = b + chunk < nblk ? b + chunk : nblk;
            T a[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = identity;
            for (; b + 7 < end; b += 8) {
                T v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = load_cg(partials + size_t(b + k) * rows + r);
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = op(a[k], v[k]);
            }
            for (; b < end; ++b) a[0] = op(a[0], load_cg(partials + size_t(b) * rows + r));
            s_acc[size_t(piece) * rows + r]
                = op(op(op(a[0], a[1]), op(a[2], a[3])), op(op(a[4], a[5]), op(a[6], a[7])));
        }
        __syncthreads();
        if (size_t(tid) < rows) {
            T a = s_acc[tid];
            for (unsigned q = 1; q < pieces; ++q) a = op(a, s_acc[size_t(q) * rows + tid]);
            out[tid] = a;
        }
    } else {
        for (size_t r = tid; r < rows; r += THREADS) {
            T a0 = identity, a1 = identity, a2 = identity, a3 = identity;
            unsigned b = 0;
            for (; b + 3 < nblk; b += 4) {
                a0 = op(a0, load_cg(partials + size_t(b) * rows + r));
                a1 = op(a1, load_cg(partials + size_t(b + 1) * rows + r));
                a2 = op(a2, load_cg(partials + size_t(b + 2) * rows + r));
                a3 = op(a3, load_cg(partials + size_t(b + 3) * rows + r));
            }
            for (; b < nblk; ++b) a0 = op(a0, load_cg(partials + size_t(b) * rows + r));
            out[r] = op(op(a0, a1), op(a2, a3));
        }
    }
    if (tid == 0) *ws.ticket = 0;
    if (!exchange) return;

    // fused cross-partition fold: publish this partition's row, the last arrival folds
    __threadfence_system();
    __syncthreads();
    if (tid == 0)
        s_last = atomicInc_system(xchg.arrivals, unsigned(xchg.nparts - 1)) == unsigned(xchg.nparts - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence_system();
    const T* all = static_cast<const T*>(xchg.rows);
    T* final_out = static_cast<T*>(xchg.final_out);
    for (size_t r = tid; r < rows; r += THREADS) {
        T a = identity;
        for (int q = 0; q < xchg.nparts; ++q) a = op(a, load_cg(all + size_t(q) * rows + r));
        final_out[r] = a;
    }
    if (xchg.done_flag != nullptr) {
        __threadfence_system();
        __syncthreads();
        if (tid == 0) *reinterpret_cast<volatile unsigned*>(xchg.done_flag) = xchg.done_ticket;
    }
}

// out[r] = fold over p (fixed order 0..nparts-1) of parts[p][r], seeded with identity
// (/root/reference scalix/algorithm/reduce_last_dim.cuh:177,208-211: result is filled
// with identity, then every device's partial row is folded in).
template<class T, class Op>
__global__ void fold_rows_kernel(
    const T* const* __restrict__ parts,  // nparts row pointers (may be peer memory)
    int nparts,
    size_t rows,
    T identity,
    T* __restrict__ out,
    Op op
) {
    for (size_t r = size_t(blockIdx.x) * blockDim.x + threadIdx.x; r < rows;
         r += size_t(gridDim.x) * blockDim.x) {
        T a = identity;
        for (int p = 0; p < nparts; ++p) a = op(a, parts[p][r]);
        out[r] = a;
    }
}

// contiguous variant: parts laid out [nparts][rows] (e.g. an NCCL all-gather buffer)
template<class T, class Op>
__global__ void fold_rows_contig_kernel(
    const T* __restrict__ parts,
    int nparts,
    size_t rows,
    T identity,
    T* __restrict__ out,
    Op op
) {
    for (size_t r = size_t(blockIdx.x) * blockDim.x + threadIdx.x; r < rows;
         r += size_t(gridDim.x) * blockDim.x) {
        T a = identity;
        for (int p = 0; p < nparts; ++p) a = op(a, parts[size_t(p) * rows + r]);
        out[r] = a;
    }
}

// *out = fold over i of map(in[i]), seeded with `identity` (count_if, min/max_element).
// Same structure as reduce_kernel: grid-stride with UNROLL independent loads per thread,
// block fold, per-CTA partials, last CTA (atomic ticket) folds them in fixed order.
template<class T, class R, class Map, class Op, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS) map_reduce_kernel(
    const T* __restrict__ in,
    size_t n,
    R identity,
    R* __restrict__ out,
    reduce_workspace_t ws,
    Map map,
    Op op
) {
    __shared__ R s_red[THREADS / 32];
    __shared__ bool s_last;
    R acc[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc[u] = identity;
    const size_t step = size_t(gridDim.x) * THREADS;
    size_t i          = size_t(blockIdx.x) * THREADS + threadIdx.x;
    for (; i + (UNROLL - 1) * step < n; i += UNROLL * step) {
        T v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = in[i + u * step];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) acc[u] = op(acc[u], map(v[u]));
    }
    for (; i < n; i += step) acc[0] = op(acc[0], map(in[i]));
#pragma unroll
    for (int u = 1; u < UNROLL; ++u) acc[0] = op(acc[0], acc[u]);
    R block = block_reduce<R, Op, THREADS>(acc[0], op, s_red);

    R* partials = static_cast<R*>(ws.partials);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = block;
        __threadfence();
        s_last = (atomicAdd(ws.ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    R a = identity;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += THREADS) a = op(a, load_cg(partials + b));
    __syncthreads();
    a = block_reduce<R, Op, THREADS>(a, op, s_red);
    if (threadIdx.x == 0) {
        *out       = a;
        *ws.ticket = 0;
    }
}

}  // namespace sclx::b200
