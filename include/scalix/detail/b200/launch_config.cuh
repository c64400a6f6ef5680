// scalix-b200: host-side launch geometry for the generic dispatcher.
//
// The reference launches ceil(n/block) CTAs per device, capped by the user's grid_size and
// traits::kernel::max_thread_count (/root/reference scalix/detail/execute_kernel.cuh:137-145).
// Here the default grid is 64 resident waves (occupancy x SM count x 64, a multiple of the
// 148 SMs): the kernel's grid-stride loop covers the rest, which removes millions of CTA
// launches for 2^30-element ranges while keeping every warp's accesses coalesced.  An explicit grid_size from the user is honoured exactly like the reference.
#pragma once

#include <cuda_runtime.h>
#include <cstddef>
#include <limits>

namespace sclx::b200 {

constexpr size_t kOversubscribe = 64;

inline int device_sm_count() {
    // per-device cache; the ordinal count of a node is small
    static int cache[64] = {0};
    int dev              = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && cache[dev] > 0) return cache[dev];
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
    if (dev >= 0 && dev < 64) cache[dev] = sms;
    return sms;
}

template<class Kernel>
inline cudaError_t range_launch_grid(
    Kernel kernel,
    size_t elements,
    unsigned block_threads,
    size_t dynamic_smem,
    size_t user_grid_size,  // SIZE_MAX = default
    unsigned* grid_out,
    size_t max_thread_count = std::numeric_limits<size_t>::max(),
    size_t leading_elements = 1  // product of the range's leading extents (1 for rank 1)
) {
    size_t blocks = (elements + block_threads - 1) / block_threads;
    if (blocks == 0) blocks = 1;
    if (user_grid_size != std::numeric_limits<size_t>::max()) {
        if (blocks > user_grid_size) blocks = user_grid_size;
    } else {
        int per_sm    = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &per_sm, kernel, int(block_threads), dynamic_smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
        // kOversubscribe resident waves, not one: the loop still strides by the whole grid
        // (coalesced on every trip), but CTAs of later waves start staggered, which
        // de-synchronises the load / store phases of the machine.  Measured on B200 with
        // 2^30-element ranges: axpy lambda 4747 -> 4942 GB/s, random-gather lambda
        // 38.4 -> 45-48 G loads/s; one CTA per 256 elements (the reference's geometry, 4 M CTAs)
        // drops to 3930 GB/s (profiles/r01_dispatch_tune.txt, profiles/r01_serve_tune.txt).
        const size_t resident = size_t(per_sm) * size_t(device_sm_count()) * kOversubscribe;
        if (blocks > resident) blocks = resident;
    }
    if (max_thread_count != std::numeric_limits<size_t>::max()) {
        size_t cap = max_thread_count / block_threads;
        if (cap < 1) cap = 1;
        if (blocks > cap) blocks = cap;
    }
    if (blocks > 0x7fffffffull) blocks = 0x7fffffffull;
    if (blocks == 0) blocks = 1;
    // Rank > 1: make the grid stride a whole number of "columns" (multiple of the leading
    // extents' product) when that costs at most a few percent of the grid: a thread's
    // successive elements then differ in the LAST coordinate only and the kernel advances its
    // md-index with one add instead of a carry chain (dispatch.cuh)
    if (leading_elements > 1 && user_grid_size == std::numeric_limits<size_t>::max()) {
        size_t a = leading_elements, b = block_threads;
        while (b) {
            const size_t r = a % b;
            a = b;
            b = r;
        }
        const size_t unit = leading_elements / a;  // grid must be a multiple of this
        if (unit > 1 && blocks >= 32 * unit) blocks -= blocks % unit;
    }
    *grid_out = unsigned(blocks);
    return cudaSuccess;
}

}  // namespace sclx::b200
