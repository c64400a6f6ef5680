// scalix-b200: sclx_b200_reduce_last_dim_noalloc — reduce_last_dim for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// No counterpart in the reference.
//
// Same kernel, same launch geometry and therefore the same bits as sclx_b200_reduce_last_dim
// (scalix_b200.cu); the only difference is where the kernel's scratch lives: the per-CTA
// partial rows and the CTA ticket sit in a file-scope `static __device__` array of fixed size
// instead of a grow-on-demand cudaMalloc workspace.  Consequences, all documented in
// include/scalix_b200.h:
//   * every GPU that loads the library holds the array (kNoAllocScratchBytes + 256 bytes)
//     until the process exits;
//   * the problem is limited: rows * sizeof(element) <= kNoAllocMaxRowBytes;
//   * ALL calls on a device share the one array and it keeps the last call's contents: the
//     ticket word is cleared on the caller's stream before every launch, the partial rows are
//     fully overwritten before they are read; two calls that overlap on the GPU (different
//     streams, no ordering between them) race.
// In host code the array's name is a host-side stand-in, not a device address: the address the
// kernel needs comes from cudaGetSymbolAddress, once per device, with that device current.
#include "scalix_b200.h"

#include "scalix/detail/b200/reduce.cuh"

#include <cstring>
#include <mutex>
#include <string>

extern "C" unsigned long long sclx_b200_note_launches(unsigned long long n);
extern "C" int sclx_b200_set_error(int code, const char* msg);

using namespace sclx::b200;

namespace {

constexpr size_t kNoAllocMaxRowBytes  = 2048;  // rows * sizeof(element): 512 f32/i32/u32 or 256 f64/i64/u64 rows
constexpr unsigned kNoAllocMaxGridX   = 1024;  // CTAs whose partial rows fit (a B200 launches 592)
constexpr size_t kNoAllocScratchBytes = size_t(kNoAllocMaxGridX) * kNoAllocMaxRowBytes;  // 2 MiB
constexpr size_t kNoAllocTicketBytes  = 256;

// [0, kNoAllocScratchBytes): partial rows; [kNoAllocScratchBytes, +256): the CTA ticket
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes + kNoAllocTicketBytes];

constexpr int kMaxDevices = 64;
struct device_slot {
    unsigned char* scratch = nullptr;  // DEVICE address of g_noalloc_scratch on this GPU
    int sm_count           = 0;
};
device_slot g_slot[kMaxDevices];
std::mutex g_slot_mutex;

int cuda_error(cudaError_t e, const char* what) {
    const std::string msg = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    sclx_b200_set_error(int(e), msg.c_str());
    return int(e);
}

// once per device, with that device current
int slot_for_current_device(device_slot** out) {
    int dev       = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_error(e, "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) return sclx_b200_set_error(SCLX_E_RUNTIME, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_slot_mutex);
    device_slot& s = g_slot[dev];
    if (s.scratch == nullptr) {
        void* address = nullptr;
        e             = cudaGetSymbolAddress(&address, g_noalloc_scratch);
        if (e != cudaSuccess) return cuda_error(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        int sms = 0;
        e       = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return cuda_error(e, "cudaDeviceGetAttribute");
        s.sm_count = sms;
        s.scratch  = static_cast<unsigned char*>(address);
    }
    *out = &s;
    return 0;
}

// the operators of scalix_b200.cu (thrust::plus / multiplies / maximum / minimum semantics)
struct na_plus {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const { return a + b; }
};
struct na_mul {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const { return a * b; }
};
struct na_max {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const { return a < b ? b : a; }
};
struct na_min {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const { return a < b ? a : b; }
};

// geometry of launch_reduce_last_dim in scalix_b200.cu (16 loads in flight, 2 accumulator sets,
// 4 CTAs/SM): the fold order depends only on (shape, grid), so results are bit-identical
constexpr int kThreads = 256, kUnroll = 16, kAccs = 2;

template<class T, class Op>
int launch(const void* in_, size_t rows, size_t cols, const void* identity_, void* out_, cudaStream_t stream, Op op) {
    device_slot* slot = nullptr;
    if (int rc = slot_for_current_device(&slot)) return rc;
    const T* in = static_cast<const T*>(in_);
    T* out      = static_cast<T*>(out_);
    T identity;
    std::memcpy(&identity, identity_, sizeof(T));
    constexpr int V    = vec_traits<T>::width;
    const bool vec_ok  = (rows % V == 0) && is_aligned16(in);
    const size_t vw    = vec_ok ? V : 1;
    const size_t slots = rows / vw;
    // rows * sizeof(T) <= 2048 bytes  =>  slots <= 512
    const size_t groups = slots <= size_t(kThreads) ? kThreads / slots : 1;
    const unsigned grid_y
        = slots <= size_t(kThreads) ? 1u : unsigned((slots + kThreads - 1) / kThreads);
    const size_t col_work = (cols + groups - 1) / groups;
    unsigned grid_x       = persistent_grid(col_work, 4, slot->sm_count);
    if (grid_y > 1) {
        unsigned per_band = (unsigned(slot->sm_count) * 4 + grid_y - 1) / grid_y;
        if (per_band < 1) per_band = 1;
        if (grid_x > per_band) grid_x = per_band;
    }
    if (size_t(grid_x) * rows * sizeof(T) > kNoAllocScratchBytes)
        return sclx_b200_set_error(
            SCLX_E_INVALID_ARGUMENT,
            "reduce_last_dim_noalloc: this GPU launches more CTAs than the static scratch (2 MiB) holds partial "
            "rows for; use sclx_b200_reduce_last_dim");
    unsigned* ticket = reinterpret_cast<unsigned*>(slot->scratch + kNoAllocScratchBytes);
    // the array keeps the previous call's contents: the ticket must start at zero (a memset
    // node allocates nothing and is capturable); the partial rows are written before they are read
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned), stream);
    if (e != cudaSuccess) return cuda_error(e, "cudaMemsetAsync(ticket)");
    reduce_workspace_t ws{slot->scratch, ticket};
    const size_t smem = size_t(kThreads) * vw * sizeof(T);
    const dim3 grid(grid_x, grid_y);
    if (vec_ok && V > 1)
        reduce_last_dim_kernel<T, Op, kThreads, kUnroll, V, kAccs>
            <<<grid, kThreads, smem, stream>>>(in, rows, cols, identity, out, ws, op);
    else
        reduce_last_dim_kernel<T, Op, kThreads, kUnroll, 1, kAccs>
            <<<grid, kThreads, smem, stream>>>(in, rows, cols, identity, out, ws, op);
    sclx_b200_note_launches(1);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_error(e, "kernel launch");
    return 0;
}

template<class T>
int launch_op(int op, const void* in, size_t rows, size_t cols, const void* identity, void* out, cudaStream_t s) {
    switch (op) {
        case SCLX_PLUS: return launch<T, na_plus>(in, rows, cols, identity, out, s, na_plus{});
        case SCLX_MULTIPLIES: return launch<T, na_mul>(in, rows, cols, identity, out, s, na_mul{});
        case SCLX_MAX: return launch<T, na_max>(in, rows, cols, identity, out, s, na_max{});
        case SCLX_MIN: return launch<T, na_min>(in, rows, cols, identity, out, s, na_min{});
        default: return sclx_b200_set_error(SCLX_E_UNSUPPORTED, "unknown operator");
    }
}

}  // namespace

extern "C" {

size_t sclx_b200_noalloc_max_row_bytes(void) { return kNoAllocMaxRowBytes; }

int sclx_b200_reduce_last_dim_noalloc(const void* in, size_t rows, size_t cols, int dtype, int op,
                                      const void* identity, void* out, void* stream) {
    if (!identity || (!out && rows) || (!in && rows * cols))
        return sclx_b200_set_error(SCLX_E_INVALID_ARGUMENT, "reduce_last_dim_noalloc: null pointer");
    const size_t es = sclx_b200_dtype_size(dtype);
    if (es == 0) return sclx_b200_set_error(SCLX_E_UNSUPPORTED, "unknown dtype");
    if (rows == 0) return 0;
    if (rows > kNoAllocMaxRowBytes / es) {
        const std::string msg = "reduce_last_dim_noalloc: rows * sizeof(element) = " + std::to_string(rows * es)
                              + " bytes exceeds the static scratch limit of " + std::to_string(kNoAllocMaxRowBytes)
                              + " bytes per row (512 4-byte or 256 8-byte rows); use sclx_b200_reduce_last_dim";
        return sclx_b200_set_error(SCLX_E_INVALID_ARGUMENT, msg.c_str());
    }
    const bool wrap = (op == SCLX_PLUS || op == SCLX_MULTIPLIES);  // integers wrap: run on the unsigned type
    auto s          = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case SCLX_I32:
            return wrap ? launch_op<uint32_t>(op, in, rows, cols, identity, out, s)
                        : launch_op<int32_t>(op, in, rows, cols, identity, out, s);
        case SCLX_U32: return launch_op<uint32_t>(op, in, rows, cols, identity, out, s);
        case SCLX_I64:
            return wrap ? launch_op<unsigned long long>(op, in, rows, cols, identity, out, s)
                        : launch_op<long long>(op, in, rows, cols, identity, out, s);
        case SCLX_U64: return launch_op<unsigned long long>(op, in, rows, cols, identity, out, s);
        case SCLX_F32: return launch_op<float>(op, in, rows, cols, identity, out, s);
        case SCLX_F64: return launch_op<double>(op, in, rows, cols, identity, out, s);
        default: return sclx_b200_set_error(SCLX_E_UNSUPPORTED, "unknown dtype");
    }
}

}  // extern "C"
