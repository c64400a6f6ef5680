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

// body of the range kernel: shared by sclx_kernel and every REGISTER_SCALIX_KERNEL_TAG kernel
template<class F, uint RangeRank, uint ThreadBlockRank>
__device__ __forceinline__ void
range_kernel_body(F& f, sclx::kernel_info<RangeRank, ThreadBlockRank>& metadata) {
    const sclx::md_range_t<RangeRank>& global_range = metadata.global_range();
    const size_t end    = metadata.start_linear() + metadata.device_range().elements();
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    size_t linear       = metadata.global_thread_linear_id();
    if (linear >= end) return;

    sclx::md_index_t<RangeRank> idx
        = sclx::md_index_t<>::create_from_linear(linear, global_range);
    const sclx::md_index_t<RangeRank> step
        = sclx::md_index_t<>::create_from_linear(stride, global_range);
    while (true) {
        f(idx, metadata);
        linear += stride;
        if (linear >= end) break;
        advance_index(idx, step, global_range);
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
sclx_kernel(F f, sclx::kernel_info<RangeRank, ThreadBlockRank> metadata) {
    range_kernel_body(f, metadata);
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
