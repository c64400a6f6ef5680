// scalix-b200: the generic per-element kernels behind kernel_handler::launch.
//
// Same contract as the two __global__ templates of
// /root/reference scalix/detail/execute_kernel.cuh:53-98 — call f(md_index, kernel_info)
// for every index of this device's contiguous sub-range (range launch), or for every
// generated index whose last coordinate falls in this device's slice (index-generator
// launch) — but:
//   * all linear arithmetic is 64-bit (no wrap at 2^32 elements per device);
//   * the md-index is advanced incrementally with carry propagation instead of being
//     rebuilt with (Rank-1) 64-bit div+mod per element per trip;
//   * the grid is sized to the machine (a fixed number of resident waves, multiples of the
//     SM count: launch_config.cuh) and the loop strides by the whole grid, so consecutive
//     threads touch consecutive elements (coalesced) on every trip.
#pragma once

#include "../../kernel_info.cuh"

namespace sclx::detail {

// idx += step (both decomposed against `shape`), propagating carries; the last
// coordinate is unbounded
template<uint Rank>
__device__ __forceinline__ void advance_index(
    md_index_t<Rank>& idx,
    const md_index_t<Rank>& step,
    const shape_t<Rank>& shape
) {
    if constexpr (Rank == 1) {
        idx[0] += step[0];
    } else {
        size_t carry = 0;
#pragma unroll
        for (uint i = 0; i + 1 < Rank; ++i) {
            size_t v = idx[i] + step[i] + carry;
            carry    = 0;
            if (v >= shape[i]) {
                v -= shape[i];
                carry = 1;
            }
            idx[i] = v;
        }
        idx[Rank - 1] += step[Rank - 1] + carry;
    }
}

// Arrays the lambda most likely walks element by element (found on the host: runtime arrays
// captured by value whose size is 1, 2, 4 or 8 bytes per range element, the results among
// them).  While trip k's loads are in flight the dispatcher pulls trip k+1's elements into L2:
// an opaque lambda whose store may alias its loads cannot overlap two elements by itself
// (2048 threads x 2 scalar loads = 16 KiB in flight per SM, 0.76 of the copy peak for axpy);
// with the next trip already on its way the same lambda reaches the rate of a hand-vectorised
// kernel (experiments/dispatch_tune.cu, profiles/r02_dispatch_tune.txt: 4941 -> 6433 GB/s).
// A prefetch never faults and never changes a result.
struct prefetch_plan_t {
    static constexpr int max_arrays = 4;
    const char* base[max_arrays];  // slots >= count repeat slot 0 (see prefetch_trip)
    unsigned elem_bytes[max_arrays];
    int count = 0;

    __host__ void add(const void* array_base, unsigned bytes_per_element) {
        if (count >= max_arrays) return;
        base[count]       = static_cast<const char*>(array_base);
        elem_bytes[count] = bytes_per_element;
        ++count;
        for (int k = count; k < max_arrays; ++k) {
            base[k]       = base[0];
            elem_bytes[k] = elem_bytes[0];
        }
    }
};

// one prefetch per slot, no per-slot predicate and no per-thread state: the address is rebuilt
// from the (uniform) plan and the next trip's linear index — a multiply-add and a prefetch per
// slot and trip; the host fills unused slots with a copy of slot 0 (a repeated prefetch of a
// line that is already on its way costs one L2 lookup)
template<int SLOTS>
__device__ __forceinline__ void prefetch_trip(const prefetch_plan_t& plan, size_t next_linear) {
#pragma unroll
    for (int k = 0; k < SLOTS; ++k)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(plan.base[k] + next_linear * plan.elem_bytes[k]));
}

// body of the range kernel: shared by sclx_kernel and every REGISTER_SCALIX_KERNEL_TAG kernel
template<class F, uint RangeRank, uint ThreadBlockRank>
__device__ __forceinline__ void range_kernel_body(
    F& f, sclx::kernel_info<RangeRank, ThreadBlockRank>& metadata, const prefetch_plan_t& plan = prefetch_plan_t{}
) {
    const sclx::md_range_t<RangeRank>& global_range = metadata.global_range();
    const size_t end    = metadata.start_linear() + metadata.device_range().elements();
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    size_t linear       = metadata.global_thread_linear_id();
    if (linear >= end) return;

    sclx::md_index_t<RangeRank> idx
        = sclx::md_index_t<>::create_from_linear(linear, global_range);
    const sclx::md_index_t<RangeRank> step
        = sclx::md_index_t<>::create_from_linear(stride, global_range);
    // the host sizes the grid so that, whenever it can, the stride is a whole number of
    // "columns": only the last coordinate moves and the carry chain is skipped (uniform branch)
    bool last_only = true;
#pragma unroll
    for (uint i = 0; i + 1 < RangeRank; ++i) last_only &= step[i] == 0;
    const int ahead = plan.count;
    while (true) {
        const size_t next = linear + stride;
        const bool more   = next < end;
        if (more && ahead > 0) {  // trip k+1 on its way to L2 while trip k runs
            if (ahead <= 2) prefetch_trip<2>(plan, next);
            else prefetch_trip<prefetch_plan_t::max_arrays>(plan, next);
        }
        f(idx, metadata);
        if (!more) break;
        linear = next;
        if (RangeRank == 1 || last_only) idx[RangeRank - 1] += step[RangeRank - 1];
        else advance_index(idx, step, global_range);
        metadata.increment_stride_count();
    }
}

// body of the index-generator kernel: every partition walks the whole generator range
// and acts only on the indices it owns
template<class F, class IndexGenerator, uint ThreadBlockRank>
__device__ __forceinline__ void generator_kernel_body(
    F& f,
    sclx::kernel_info<IndexGenerator::range_rank, ThreadBlockRank>& metadata,
    IndexGenerator& idx_gen,
    sclx::index_t slice_idx,
    sclx::index_t slice_size
) {
    constexpr uint range_rank = IndexGenerator::range_rank;
    constexpr uint index_rank = IndexGenerator::index_rank;
    const sclx::md_range_t<range_rank>& global_range = metadata.global_range();
    const size_t end    = global_range.elements();
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    size_t linear       = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (linear >= end) return;

    sclx::md_index_t<range_rank> gen_idx
        = sclx::md_index_t<>::create_from_linear(linear, global_range);
    const sclx::md_index_t<range_rank> step
        = sclx::md_index_t<>::create_from_linear(stride, global_range);
    while (true) {
        const sclx::md_index_t<index_rank> idx = idx_gen(gen_idx);
        // owner filter: only the partition holding idx's last coordinate acts on it
        if (idx[index_rank - 1] - slice_idx < slice_size) f(idx, metadata);
        linear += stride;
        if (linear >= end) break;
        advance_index(gen_idx, step, global_range);
        metadata.increment_stride_count();
    }
}

template<class F, uint RangeRank, uint ThreadBlockRank>
__global__ void
sclx_kernel(F f, sclx::kernel_info<RangeRank, ThreadBlockRank> metadata, prefetch_plan_t plan) {
    range_kernel_body(f, metadata, plan);
}

template<class F, class IndexGenerator, uint ThreadBlockRank>
__global__ void sclx_kernel(
    F f,
    sclx::kernel_info<IndexGenerator::range_rank, ThreadBlockRank> metadata,
    IndexGenerator idx_gen,
    sclx::index_t slice_idx,
    sclx::index_t slice_size
) {
    generator_kernel_body(f, metadata, idx_gen, slice_idx, slice_size);
}

}  // namespace sclx::detail
