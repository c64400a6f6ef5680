/* scalix-b200 — C ABI of the B200-native Scalix hot path.
 *
 * The reference (NAGAGroup/Scalix) is a header-only C++17/CUDA template library
 * with no FFI of its own: its boundary is the `sclx::` header API.  This library
 * sits directly UNDER that header API (include/scalix/ in this repository binds to
 * it) and is what a foreign-language binding (ctypes, cgo, JNI ...) of the hot
 * path would bind.  Every entry point names the reference interface it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; `stream` is a cudaStream_t passed as void*
 *     (NULL = the legacy default stream);
 *   - device pointers unless the name ends in `_host`;
 *   - return value: 0 on success, a positive cudaError_t on a CUDA failure, a
 *     negative SCLX_E_* code on invalid arguments; sclx_b200_last_error() gives
 *     the message (thread-local).  Argument errors correspond to the
 *     std::invalid_argument / std::runtime_error throws of the reference
 *     (/root/reference scalix/algorithm/inclusive_scan.cuh:52-55,
 *      scalix/algorithm/reduce_last_dim.cuh:171-174);
 *   - all kernels are asynchronous on `stream`; nothing here synchronises unless
 *     documented.  There is NO CPU fallback: without a CUDA device every compute
 *     entry point fails with the CUDA error.
 *   - memory layout is the reference's: first index fastest, last dimension
 *     slowest and partitioned (/root/reference scalix/index.cuh:53-64).
 */
#ifndef SCALIX_B200_H
#define SCALIX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCLX_B200_ABI_VERSION 1

/* element types of the framework-owned kernels */
typedef enum sclx_dtype {
    SCLX_I32 = 0,
    SCLX_U32 = 1,
    SCLX_I64 = 2,
    SCLX_U64 = 3,
    SCLX_F32 = 4,
    SCLX_F64 = 5
} sclx_dtype;

/* binary operators: the aliases of /root/reference scalix/algorithm/functional.cuh:38-60
 * (plus, multiplies, maximum/max, minimum/min).  Integer plus/multiplies wrap. */
typedef enum sclx_op {
    SCLX_PLUS       = 0,
    SCLX_MULTIPLIES = 1,
    SCLX_MAX        = 2,
    SCLX_MIN        = 3
} sclx_op;

enum {
    SCLX_E_INVALID_ARGUMENT = -1, /* std::invalid_argument in the reference   */
    SCLX_E_UNSUPPORTED      = -2, /* dtype/op combination not instantiated     */
    SCLX_E_RUNTIME          = -3  /* std::runtime_error in the reference       */
};

/* ---- library ---------------------------------------------------------------- */
int sclx_b200_abi_version(void);
const char* sclx_b200_last_error(void);
/* replaces sclx::cuda::traits::device_count() (/root/reference scalix/cuda.hpp:100-112) */
int sclx_b200_device_count(void);
/* replaces sclx::cuda::set_device (/root/reference scalix/cuda.hpp:61-70) */
int sclx_b200_set_device(int device);
size_t sclx_b200_dtype_size(int dtype);

/* ---- partition arithmetic (host only, no CUDA) --------------------------------
 * replaces the split computation of array::set_primary_devices
 * (/root/reference scalix/array.cuh:467-513): ceil(last_dim/ndev) columns per
 * device, surplus devices dropped.  Writes up to `ndev` (start, len) pairs and
 * returns the number of devices actually used (>= 1), or a negative error. */
int sclx_b200_partition(size_t last_dim, int ndev, size_t* start, size_t* len);

/* ---- memory --------------------------------------------------------------------
 * replaces detail::allocate_cuda_usm (/root/reference scalix/array.cuh:241-248) for
 * one partition: explicit device memory on the current device, zero-filled (the
 * reference's examples rely on fresh managed memory reading as zero,
 * examples/distributed_indexing.cu:79,146-154). */
int sclx_b200_malloc(void** ptr, size_t bytes);
int sclx_b200_free(void* ptr);
/* pinned host staging memory for the *_host entry points */
int sclx_b200_host_alloc(void** ptr, size_t bytes);
int sclx_b200_host_free(void* ptr);
int sclx_b200_memcpy_h2d(void* dst, const void* src, size_t bytes, void* stream);
int sclx_b200_memcpy_d2h(void* dst, const void* src, size_t bytes, void* stream);
int sclx_b200_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream);
int sclx_b200_stream_create(void** stream);
int sclx_b200_stream_destroy(void* stream);
int sclx_b200_stream_synchronize(void* stream);
/* cross-process peer mapping of a partition (multi-process ranks): 64-byte handles */
int sclx_b200_ipc_get_handle(void* ptr, void* handle64);
int sclx_b200_ipc_open_handle(const void* handle64, void** ptr);
int sclx_b200_ipc_close(void* ptr);

/* ---- kernels --------------------------------------------------------------------- */

/* sclx::fill (/root/reference scalix/fill.cuh:40-51): out[i] = *value for n elements of
 * elem_bytes in {1,2,4,8,16}. */
int sclx_b200_fill(void* out, size_t n, int elem_bytes, const void* value, void* stream);

/* sclx::iota (/root/reference scalix/iota.cuh:40-51): out[i] = start + (first_index + i).
 * `first_index` is the partition's offset in the whole array. */
int sclx_b200_iota(void* out, size_t n, int dtype, const void* start,
                   size_t first_index, void* stream);

/* sclx::algorithm::inclusive_scan (/root/reference scalix/algorithm/inclusive_scan.cuh:44-207)
 * for ONE partition: out[i] = in[0] (op) ... (op) in[i] over flattened storage order, in the
 * element type (integers wrap).  `lower_totals` (device memory, n_lower values, may be NULL
 * when n_lower == 0) are the un-carried totals of the partitions below this one; they are
 * chained and applied exactly as ibid. :105-157.  `total_out` (device, optional) receives
 * this partition's last output.  `in` may equal `out`. */
int sclx_b200_inclusive_scan(const void* in, void* out, size_t n, int dtype, int op,
                             const void* lower_totals, int n_lower, void* total_out,
                             void* stream);

/* sclx::algorithm::reduce (/root/reference scalix/algorithm/reduce.cuh:44-101) for one
 * partition: *out = identity (op) in[0] (op) ... ; `out` is device memory. */
int sclx_b200_reduce(const void* in, size_t n, int dtype, int op, const void* identity,
                     void* out, void* stream);

/* sclx::algorithm::reduce_last_dim (/root/reference scalix/algorithm/reduce_last_dim.cuh:141-269)
 * for one partition holding `cols` columns of `rows` contiguous elements
 * (rows = product of the leading dims): out[s] = identity (op) fold_j in[s + rows*j]. */
int sclx_b200_reduce_last_dim(const void* in, size_t rows, size_t cols, int dtype, int op,
                              const void* identity, void* out, void* stream);

/* the cross-partition fold of ibid. :208-211 in FIXED partition order:
 * out[s] = identity (op) parts[0][s] (op) parts[1][s] ...; parts is [nparts][rows]
 * contiguous device memory (e.g. an all-gather buffer). */
int sclx_b200_fold_rows(const void* parts, int nparts, size_t rows, int dtype, int op,
                        const void* identity, void* out, void* stream);

/* cross-partition gather, the access pattern of /root/reference
 * examples/scaling_performance.cu:59-78: out[i] = src[idx[i]], idx_bytes in {4,8},
 * elem_bytes in {4,8}.  src_parts: nparts device/peer pointers (host array of pointers),
 * partition p holding elements [p*elems_per_part, (p+1)*elems_per_part). */
int sclx_b200_gather(const void* idx, size_t n, int idx_bytes, const void* const* src_parts,
                     int nparts, size_t elems_per_part, int elem_bytes, void* out,
                     void* stream);

/* ---- sclx::execute_kernel range launches with fixed device lambdas ------------------------
 * These run THROUGH the generic per-element dispatcher (include/scalix/detail/b200/dispatch.cuh,
 * replacing /root/reference scalix/detail/execute_kernel.cuh:53-69,100-200), i.e. what a user's
 * kernel_handler::launch(range, result, lambda) executes, for the two lambdas BASELINE.json names.
 * `nparts` >= 1 splits the range's last dimension like set_primary_devices and issues one launch
 * per partition on the CURRENT device (logical-device emulation, the equivalent of
 * SCALIX_EMULATE_MULTIDEVICE, /root/reference CMakeLists.txt:31-44). */

/* config 3 (examples/scaling_performance.cu style): y[idx] = a*x[idx] + y[idx] over a rank-1
 * range of n elements */
int sclx_b200_axpy_f32(float a, const void* x, void* y, size_t n, int nparts, void* stream);
/* same lambda over a rank-3 range {d0, d1, last} (exercises md-index arithmetic) */
int sclx_b200_axpy_f32_rank3(float a, const void* x, void* y, size_t d0, size_t d1, size_t last,
                             int nparts, void* stream);
/* config 4 through the range launcher: out[idx] = src[index[idx]] (u64 indices, f32 payload) */
int sclx_b200_gather_launch_f32(const void* index, size_t n, const void* src, void* out,
                                int nparts, void* stream);

/* ---- host-buffer entry points (input and output in HOST memory; copies included) ------
 * Blocking.  Input is streamed to the device in chunks that overlap with the kernels. */
int sclx_b200_reduce_last_dim_host(const void* host_in, size_t rows, size_t cols, int dtype,
                                   int op, const void* identity, void* host_out);
int sclx_b200_inclusive_scan_host(const void* host_in, void* host_out, size_t n, int dtype,
                                  int op);

/* ---- allocation-free variant -----------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same kernel, launch geometry and result bits as sclx_b200_reduce_last_dim,
 * but the kernel's scratch (per-CTA partial rows + ticket) is a file-scope static __device__
 * array of 2 MiB + 256 bytes (scalix_b200/csrc/no_alloc.cu) instead of a cudaMalloc'd workspace:
 * the call allocates and frees nothing (it issues one 4-byte cudaMemsetAsync and one kernel on
 * `stream`).  Every GPU that loads the library holds the array until the process exits.
 *   - Limit: rows * sizeof(element) <= sclx_b200_noalloc_max_row_bytes() = 2048 bytes (512
 *     4-byte or 256 8-byte rows); larger problems are rejected with SCLX_E_INVALID_ARGUMENT and a
 *     message naming the limit.  `cols` is unlimited.
 *   - All calls on a device SHARE the one array and it keeps the last call's contents (the call
 *     clears the ticket word itself): calls that overlap on the GPU — different streams with no
 *     ordering between them — race.  Calls on one stream, or ordered by events, are fine.
 *   - The first call on a device resolves the array's device address (cudaGetSymbolAddress);
 *     make one warm-up call per device before starting a graph capture. */
size_t sclx_b200_noalloc_max_row_bytes(void);
int sclx_b200_reduce_last_dim_noalloc(const void* in, size_t rows, size_t cols, int dtype, int op,
                                      const void* identity, void* out, void* stream);

/* ---- runtime under the sclx:: header API (include/scalix/) --------------------------------------
 * Logical devices are the physical GPUs, or N emulated devices on GPU 0 when
 * SCLX_B200_EMULATE_DEVICES=N / sclx_b200_rt_set_emulated_devices(N) is set before first use
 * (the equivalent of SCALIX_EMULATE_MULTIDEVICE, /root/reference CMakeLists.txt:31-44). */
int sclx_b200_rt_init(void);
int sclx_b200_rt_set_emulated_devices(int n);
int sclx_b200_rt_device_count(void);              /* logical devices */
int sclx_b200_rt_physical_device(int logical);
void* sclx_b200_rt_stream(int logical);           /* the device's persistent stream (cudaStream_t) */
int sclx_b200_rt_sm_count(int logical);

/* One array = one contiguous virtual range whose pages are explicit physical allocations on
 * the partitions' devices, mapped read/write into every device (peer loads/stores over NVLink)
 * and shadowed by an on-demand host mirror at the same addresses.  Replaces
 * cudaMallocManaged + cudaMemAdvise + cudaMemPrefetchAsync of
 * /root/reference scalix/array.cuh:241-248,517-615.  Partition p starts at byte_starts[p]
 * (byte_starts[0] == 0) and lives on logical device devices[p].  Memory is zero-filled. */
int sclx_b200_array_create(size_t bytes, int nparts, const int* devices, const size_t* byte_starts,
                           void** base);
int sclx_b200_array_free(void* base);
/* array::set_primary_devices on a live array: moves the pages, keeps pointer and contents */
int sclx_b200_array_repartition(void* base, int nparts, const int* devices,
                                const size_t* byte_starts);
int sclx_b200_array_owner(const void* ptr);       /* logical device holding the page, or -1 */
/* runtime arrays whose base address appears as an 8-byte word inside `closure` (a device lambda's
 * by-value captures): up to max_out (base, bytes) pairs; returns how many were found */
int sclx_b200_rt_find_captured_arrays(const void* closure, size_t closure_bytes, int max_out,
                                      void** bases, size_t* bytes);

/* host/device coherence of the mirror */
int sclx_b200_rt_before_device_access(void);      /* write back host stores, drop host copies */
/* the range becomes host-resident: readable and writable without faults (syscalls into it
 * work), pinned for PCIe-rate copies when it covers >= 1/4 of the array, written back as a
 * whole before the next device access */
int sclx_b200_rt_prefetch_to_host(const void* ptr, size_t bytes);
int sclx_b200_rt_quiesce(void);                   /* wait for all device work */
unsigned long long sclx_b200_rt_host_faults(void);

/* launch groups: replaces the stream-per-launch + synchronise-all of
 * /root/reference scalix/detail/execute_kernel.cuh:110-115,197-199 */
int sclx_b200_rt_begin_launch(int ndev, const int* devices);
int sclx_b200_rt_end_launch(int ndev, const int* devices);
int sclx_b200_rt_stream_wait(int waiter, int signaller);
/* record `logical`'s event on its stream / make `waiter`'s stream wait for `signaller`'s event */
int sclx_b200_rt_record_done(int logical);
int sclx_b200_rt_wait_done(int waiter, int signaller);
/* queue `logical`'s next completion ticket (a 32-bit stream write into pinned host memory)
 * behind everything issued on its stream so far: sync_devices polls these words instead of
 * calling cudaStreamSynchronize per device.  Every launch group ends with it. */
int sclx_b200_rt_signal_done(int logical);
int sclx_b200_rt_sync_devices(int ndev, const int* devices);
int sclx_b200_rt_devices_idle(int ndev, const int* devices); /* 1 = nothing running (never blocks) */
/* host-polled completion: a word of pinned, mapped host memory + a process-wide ticket; the
 * kernel that finishes a multi-device operation stores the ticket, the caller polls for it
 * instead of synchronising N streams (errors fall back to a real synchronisation) */
int sclx_b200_rt_host_flag_acquire(unsigned** flag, unsigned* ticket);
int sclx_b200_rt_host_flag_release(unsigned* flag);
int sclx_b200_rt_host_flag_wait(const unsigned* flag, unsigned ticket, int ndev, const int* devices);

/* Workspaces of the scan / reduce kernels, LEASED PER CALL: one lease per (algorithm call,
 * logical device), used in stream order on that device's stream and returned when the call has
 * synchronised.  Two calls in flight on different host threads never share partials, tickets
 * or look-back descriptors (the reference keeps these function-local,
 * /root/reference scalix/algorithm/reduce_last_dim.cuh:180-211, inclusive_scan.cuh:75-123). */
int sclx_b200_rt_workspace_acquire(int logical, void** lease);
int sclx_b200_rt_workspace_release(void* lease, int poisoned); /* poisoned: re-zero tickets first */
int sclx_b200_rt_workspace_reduce(void* lease, size_t bytes, void** partials, void** ticket);
/* look-back descriptors: word_bytes 8 (elements <= 4 bytes) or 16 (8-byte elements); each
 * word size has its own buffer and epoch */
int sclx_b200_rt_workspace_scan(void* lease, size_t tiles, int word_bytes, void** desc,
                                unsigned* epoch, void** counters);
/* small peer-visible scratch (per-partition totals, partial rows), leased per call */
int sclx_b200_rt_scratch_acquire(size_t bytes, void** ptr);
int sclx_b200_rt_scratch_release(void* ptr, int poisoned);
/* Freed arrays stay mapped in a bounded pool (default min(32 GiB, device memory / 8);
 * SCLX_B200_POOL_BYTES overrides, 0 = off; sclx_b200_rt_pool_release() empties it, e.g. before a
 * large cudaMalloc of another library; sclx::cuda::query_memory_status counts it as free) and
 * are reused by the next array_create with the same size and placement: building a multi-GiB
 * peer-mapped range costs tens of milliseconds on a multi-GPU node, reusing one costs its
 * zero-fill.  (The reference pays cudaMallocManaged + 4*ndev+2 cudaMemAdvise + ndev prefetches
 * per array, scalix/array.cuh:345,522-612.) */
size_t sclx_b200_rt_pool_bytes(void);
size_t sclx_b200_rt_pool_bytes_on(int logical);  /* parked bytes on that device's GPU */
unsigned long long sclx_b200_rt_pool_hits(void);
int sclx_b200_rt_pool_release(void);
/* per-device peer-visible grow-only workspace of sclx::algorithm::gather (exchange form) */
int sclx_b200_rt_exchange_workspace(int logical, size_t bytes, void** ptr);

/* number of kernels launched in this process by this library and by the header API's launch
 * groups, which report theirs through note_launches (monotonic) */
unsigned long long sclx_b200_launch_count(void);
unsigned long long sclx_b200_note_launches(unsigned long long n);

#ifdef __cplusplus
}
#endif
#endif /* SCALIX_B200_H */
