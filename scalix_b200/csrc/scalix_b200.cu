// scalix-b200: C-ABI implementation (see include/scalix_b200.h).
// Instantiates the kernel templates of include/scalix/detail/b200/ for the standard
// element types and the reference's functional aliases and exposes them with plain
// pointers and sizes.  No CPU fallback: every compute entry point needs a CUDA device.
#include "scalix_b200.h"

#include "scalix/detail/b200/reduce.cuh"
#include "scalix/detail/b200/scan_stream.cuh"
#include "scalix/detail/b200/stream_kernels.cuh"

#include <atomic>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

using namespace sclx::b200;
using u64_t = unsigned long long;
using i64_t = long long;

namespace {

thread_local std::string g_last_error;
std::atomic<unsigned long long> g_launches{0};

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

int cuda_fail(cudaError_t e, const char* what) {
    g_last_error = std::string(what) + ": " + cudaGetErrorName(e) + " ("
                 + cudaGetErrorString(e) + ")";
    return int(e);
}

#define SCLX_CUDA(call)                                                        \
    do {                                                                       \
        cudaError_t e__ = (call);                                              \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);                  \
    } while (0)

#define SCLX_LAUNCHED()                                                        \
    do {                                                                       \
        g_launches.fetch_add(1, std::memory_order_relaxed);                    \
        cudaError_t e__ = cudaGetLastError();                                  \
        if (e__ != cudaSuccess) return cuda_fail(e__, "kernel launch");        \
    } while (0)

// ---- operators ----------------------------------------------------------------
struct op_plus {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const {
        return a + b;
    }
};
struct op_mul {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const {
        return a * b;
    }
};
struct op_max {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const {
        return a < b ? b : a;  // thrust::maximum: (lhs < rhs) ? rhs : lhs
    }
};
struct op_min {
    template<class T>
    __host__ __device__ T operator()(const T& a, const T& b) const {
        return a < b ? a : b;  // thrust::minimum: (lhs < rhs) ? lhs : rhs
    }
};

// ---- workspaces: one per (device, stream) ----------------------------------------------
// The C ABI accepts arbitrary caller streams.  Kernels on one stream use that stream's
// workspace in stream order; two streams never share partials, tickets or look-back
// descriptors, so concurrent calls on different streams neither race nor serialise.
struct stream_workspace {
    void* partials     = nullptr;
    size_t part_bytes  = 0;
    unsigned* ticket   = nullptr;  // 256 zeroed bytes: [0] reduce ticket, [32..] scan tile tickets
    // look-back descriptors: one buffer and epoch per word size (8-byte words for elements of
    // <= 4 bytes, 16-byte words for 8-byte elements) — a stale word of the other layout must
    // never be decoded as {epoch|state, value}
    void* desc[2]        = {nullptr, nullptr};
    size_t desc_bytes[2] = {0, 0};
    unsigned epoch[2]    = {0, 0};
};

struct device_state {
    std::mutex mu;
    int sm_count = 0;
    std::map<cudaStream_t, stream_workspace> streams;
};

constexpr int kMaxDevices        = 64;
constexpr size_t kMaxStreamSlots = 256;  // workspaces kept per device before a reset
device_state g_dev[kMaxDevices];

int current_state(device_state** out) {
    int dev = 0;
    SCLX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices)
        return fail(SCLX_E_RUNTIME, "device ordinal out of range");
    device_state& st = g_dev[dev];
    if (st.sm_count == 0) {
        int sms = 0;
        SCLX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        st.sm_count = sms;
    }
    *out = &st;
    return 0;
}

// the workspace of `stream` on the current device (call with st.mu held)
int workspace_for(device_state& st, cudaStream_t stream, stream_workspace** out) {
    auto it = st.streams.find(stream);
    if (it == st.streams.end()) {
        if (st.streams.size() >= kMaxStreamSlots) {
            // streams come and go: start over once nothing can still be using the old slots
            SCLX_CUDA(cudaDeviceSynchronize());
            for (auto& kv : st.streams) {
                cudaFree(kv.second.partials);
                cudaFree(kv.second.ticket);
                cudaFree(kv.second.desc[0]);
                cudaFree(kv.second.desc[1]);
            }
            st.streams.clear();
        }
        stream_workspace ws;
        SCLX_CUDA(cudaMalloc(&ws.ticket, 256));
        SCLX_CUDA(cudaMemsetAsync(ws.ticket, 0, 256, stream));
        it = st.streams.emplace(stream, ws).first;
    }
    *out = &it->second;
    return 0;
}

int ensure_bytes(void** p, size_t* have, size_t want, bool zero, cudaStream_t stream) {
    if (*have >= want) return 0;
    if (*p) {
        SCLX_CUDA(cudaStreamSynchronize(stream));  // only this stream's kernels use the buffer
        SCLX_CUDA(cudaFree(*p));
        *p    = nullptr;
        *have = 0;
    }
    size_t cap = want + want / 2 + 4096;
    SCLX_CUDA(cudaMalloc(p, cap));
    if (zero) SCLX_CUDA(cudaMemsetAsync(*p, 0, cap, stream));
    *have = cap;
    return 0;
}

// ---- dtype dispatch ----------------------------------------------------------------
size_t dtype_size(int dtype) {
    switch (dtype) {
        case SCLX_I32:
        case SCLX_U32:
        case SCLX_F32: return 4;
        case SCLX_I64:
        case SCLX_U64:
        case SCLX_F64: return 8;
        default: return 0;
    }
}

// calls f(T{}, Op{}) for the (dtype, op) pair; integer plus/multiplies run on the
// unsigned type of the same width (two's-complement wrap, bit-identical, no UB)
template<class F>
int dispatch(int dtype, int op, F&& f) {
    const bool wrap_op = (op == SCLX_PLUS || op == SCLX_MULTIPLIES);
    if (wrap_op) {
        if (dtype == SCLX_I32) dtype = SCLX_U32;
        if (dtype == SCLX_I64) dtype = SCLX_U64;
    }
#define SCLX_OP_SWITCH(T)                                                      \
    switch (op) {                                                              \
        case SCLX_PLUS: return f(T{}, op_plus{});                              \
        case SCLX_MULTIPLIES: return f(T{}, op_mul{});                         \
        case SCLX_MAX: return f(T{}, op_max{});                                \
        case SCLX_MIN: return f(T{}, op_min{});                                \
        default: return fail(SCLX_E_UNSUPPORTED, "unknown operator");          \
    }
    switch (dtype) {
        case SCLX_I32:
            switch (op) {
                case SCLX_MAX: return f(int32_t{}, op_max{});
                case SCLX_MIN: return f(int32_t{}, op_min{});
                default: return fail(SCLX_E_UNSUPPORTED, "unknown operator");
            }
        case SCLX_I64:
            switch (op) {
                case SCLX_MAX: return f(i64_t{}, op_max{});
                case SCLX_MIN: return f(i64_t{}, op_min{});
                default: return fail(SCLX_E_UNSUPPORTED, "unknown operator");
            }
        case SCLX_U32: SCLX_OP_SWITCH(uint32_t)
        case SCLX_U64: SCLX_OP_SWITCH(u64_t)
        case SCLX_F32: SCLX_OP_SWITCH(float)
        case SCLX_F64: SCLX_OP_SWITCH(double)
        default: return fail(SCLX_E_UNSUPPORTED, "unknown dtype");
    }
#undef SCLX_OP_SWITCH
}

template<class T>
T load_scalar(const void* p) {
    T v;
    std::memcpy(&v, p, sizeof(T));
    return v;
}

// ---- launch helpers -----------------------------------------------------------------
// Scan geometry (scan_stream_geometry, scan_stream.cuh): the streaming kernel for 16-byte
// aligned slices, the register look-back kernel for unaligned ones.
constexpr int kScanThreads = 256;  // register kernel
constexpr int kScanVecs    = 6;

template<class T, class Op>
int launch_scan(const T* in, T* out, size_t n, const T* lower, int n_lower, T* total_out,
                cudaStream_t stream, Op op) {
    using desc_io = tile_desc<T>;
    using geo     = scan_stream_geometry<T>;
    using scfg    = typename geo::config;
    using rcfg    = scan_config<T, kScanThreads, kScanVecs>;
    if (n == 0) return 0;
    device_state* st = nullptr;
    if (int rc = current_state(&st)) return rc;
    std::lock_guard<std::mutex> lock(st->mu);

    const bool aligned     = is_aligned16(in) && is_aligned16(out);
    const size_t tile_items = aligned ? scfg::tile_items : rcfg::tile_items;
    const size_t tiles      = (n + tile_items - 1) / tile_items;
    if (tiles > 0x7fffffffull) return fail(SCLX_E_INVALID_ARGUMENT, "scan: partition too large");
    stream_workspace* ws = nullptr;
    if (int rc = workspace_for(*st, stream, &ws)) return rc;
    constexpr int k  = desc_io::bytes == 8 ? 0 : 1;
    const size_t had = ws->desc_bytes[k];
    if (int rc = ensure_bytes(&ws->desc[k], &ws->desc_bytes[k], tiles * desc_io::bytes, true, stream))
        return rc;
    if (ws->desc_bytes[k] != had) ws->epoch[k] = 0;
    if (++ws->epoch[k] >= (1u << 30)) {
        SCLX_CUDA(cudaMemsetAsync(ws->desc[k], 0, ws->desc_bytes[k], stream));
        ws->epoch[k] = 1;
    }
    scan_carry_t carry{lower, n_lower};
    auto* desc = static_cast<typename desc_io::word_t*>(ws->desc[k]);
    if (aligned) {
        auto kernel = inclusive_scan_stream_kernel<T, Op, geo::vecs, geo::chunks, geo::stages,
                                                   geo::w_red, geo::w_scan, geo::lb_k>;
        static bool configured[kMaxDevices] = {};
        int dev = 0;
        SCLX_CUDA(cudaGetDevice(&dev));
        if (!configured[dev]) {
            SCLX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           int(scfg::smem_bytes)));
            configured[dev] = true;
        }
        // one persistent CTA per SM (the ring takes the whole shared memory)
        const unsigned grid = unsigned(tiles < size_t(st->sm_count) ? tiles : size_t(st->sm_count));
        kernel<<<grid, scfg::threads, scfg::smem_bytes, stream>>>(
            in, out, n, desc, ws->epoch[k], carry, total_out,
            reinterpret_cast<scan_stream_counters*>(ws->ticket + 32), op);
    } else {
        inclusive_scan_kernel<T, Op, kScanThreads, kScanVecs, false>
            <<<unsigned(tiles), kScanThreads, 0, stream>>>(
                in, out, n, desc, ws->epoch[k], carry, total_out, op);
    }
    SCLX_LAUNCHED();
    return 0;
}

constexpr int kReduceThreads = 256;
constexpr int kReduceVecs    = 8;

template<class T, class Op>
int launch_reduce(const T* in, size_t n, T identity, T* out, cudaStream_t stream, Op op) {
    device_state* st = nullptr;
    if (int rc = current_state(&st)) return rc;
    std::lock_guard<std::mutex> lock(st->mu);
    constexpr int V  = vec_traits<T>::width;
    const size_t work = (n / V + kReduceThreads - 1) / kReduceThreads;
    const unsigned grid = persistent_grid(work, 8, st->sm_count);
    stream_workspace* sw = nullptr;
    if (int rc = workspace_for(*st, stream, &sw)) return rc;
    if (int rc = ensure_bytes(&sw->partials, &sw->part_bytes, size_t(grid) * sizeof(T), false,
                              stream))
        return rc;
    reduce_workspace_t ws{sw->partials, sw->ticket};
    if (is_aligned16(in))
        reduce_kernel<T, Op, kReduceThreads, kReduceVecs, true>
            <<<grid, kReduceThreads, 0, stream>>>(in, n, identity, out, ws, op);
    else
        reduce_kernel<T, Op, kReduceThreads, kReduceVecs, false>
            <<<grid, kReduceThreads, 0, stream>>>(in, n, identity, out, ws, op);
    SCLX_LAUNCHED();
    return 0;
}

// geometry: profiles/r01_rld_tune.txt (16 loads in flight, 2 accumulator sets, 4 CTAs/SM)
constexpr int kRldThreads = 256;
constexpr int kRldUnroll  = 16;
constexpr int kRldAccs    = 2;

template<class T, class Op>
int launch_reduce_last_dim(const T* in, size_t rows, size_t cols, T identity, T* out,
                           cudaStream_t stream, Op op) {
    if (rows == 0) return 0;
    device_state* st = nullptr;
    if (int rc = current_state(&st)) return rc;
    std::lock_guard<std::mutex> lock(st->mu);
    constexpr int V   = vec_traits<T>::width;
    const bool vec_ok = (rows % V == 0) && is_aligned16(in);
    const size_t vw    = vec_ok ? V : 1;
    const size_t slots = rows / vw;
    size_t groups      = slots <= size_t(kRldThreads) ? kRldThreads / slots : 1;
    unsigned grid_y    = slots <= size_t(kRldThreads)
                           ? 1u
                           : unsigned((slots + kRldThreads - 1) / kRldThreads);
    if (grid_y > 65535u)
        return fail(SCLX_E_UNSUPPORTED, "reduce_last_dim: leading extent too large");
    size_t col_work = (cols + groups - 1) / groups;
    // enough CTAs to fill the machine; each keeps kRldUnroll vectors/thread in flight
    unsigned ctas_per_sm = 4;
    unsigned grid_x = persistent_grid(col_work, ctas_per_sm, st->sm_count);
    if (grid_y > 1) {
        unsigned per_band = (unsigned(st->sm_count) * ctas_per_sm + grid_y - 1) / grid_y;
        if (per_band < 1) per_band = 1;
        if (grid_x > per_band) grid_x = per_band;
    }
    stream_workspace* sw = nullptr;
    if (int rc = workspace_for(*st, stream, &sw)) return rc;
    if (int rc = ensure_bytes(&sw->partials, &sw->part_bytes,
                              size_t(grid_x) * rows * sizeof(T), false, stream))
        return rc;
    reduce_workspace_t ws{sw->partials, sw->ticket};
    const size_t smem = size_t(kRldThreads) * vw * sizeof(T);
    dim3 grid(grid_x, grid_y);
    if (vec_ok && V > 1)
        reduce_last_dim_kernel<T, Op, kRldThreads, kRldUnroll, V, kRldAccs>
            <<<grid, kRldThreads, smem, stream>>>(in, rows, cols, identity, out, ws, op);
    else
        reduce_last_dim_kernel<T, Op, kRldThreads, kRldUnroll, 1, kRldAccs>
            <<<grid, kRldThreads, smem, stream>>>(in, rows, cols, identity, out, ws, op);
    SCLX_LAUNCHED();
    return 0;
}

template<class T>
int launch_fill(T* out, size_t n, T value, cudaStream_t stream) {
    if (n == 0) return 0;
    device_state* st = nullptr;
    if (int rc = current_state(&st)) return rc;
    constexpr int V   = vec_traits<T>::width;
    const size_t work = (n / V + 1023) / 1024;
    const unsigned grid = persistent_grid(work, 8, st->sm_count, kStreamWaves);
    fill_kernel<T, 256, 4><<<grid, 256, 0, stream>>>(out, n, value);
    SCLX_LAUNCHED();
    return 0;
}

template<class T>
int launch_iota(T* out, size_t n, T start, size_t first_index, cudaStream_t stream) {
    if (n == 0) return 0;
    device_state* st = nullptr;
    if (int rc = current_state(&st)) return rc;
    constexpr int V   = vec_traits<T>::width;
    const size_t work = (n / V + 1023) / 1024;
    const unsigned grid = persistent_grid(work, 8, st->sm_count, kStreamWaves);
    iota_kernel<T, 256, 4><<<grid, 256, 0, stream>>>(out, n, start, first_index);
    SCLX_LAUNCHED();
    return 0;
}

template<class T, class I>
int launch_gather(const I* idx, size_t n, const gather_parts_t& parts, T* out,
                  cudaStream_t stream) {
    if (n == 0) return 0;
    device_state* st = nullptr;
    if (int rc = current_state(&st)) return rc;
    // NOT a persistent grid: a random-access kernel runs ~30 % faster when many small CTAs (one
    // 1024-index chunk each) start staggered than when one resident wave sweeps in lock step
    // (experiments/serve_tune.cu, profiles/r01_serve_tune.txt)
    (void)st;
    const size_t work   = (n + 1023) / 1024;
    const unsigned grid = unsigned(work < 0x7fffffffull ? work : 0x7fffffffull);
    gather_kernel<T, I, 256, 4><<<grid, 256, 0, stream>>>(idx, n, parts, out);
    SCLX_LAUNCHED();
    return 0;
}

// NOTE: these are named functors, not generic lambdas, on purpose: nvcc replaces a kernel
// launch whose instantiation it first meets inside a generic lambda body by exit(1).
struct scan_call {
    const void* in; void* out; size_t n; const void* lower; int n_lower; void* total_out;
    cudaStream_t stream;
    template<class T, class Op>
    int operator()(T, Op o) const {
        return launch_scan<T, Op>((const T*)in, (T*)out, n, (const T*)lower, n_lower, (T*)total_out,
                                  stream, o);
    }
};
struct reduce_call {
    const void* in; size_t n; const void* identity; void* out; cudaStream_t stream;
    template<class T, class Op>
    int operator()(T, Op o) const {
        return launch_reduce<T, Op>((const T*)in, n, load_scalar<T>(identity), (T*)out, stream, o);
    }
};
struct rld_call {
    const void* in; size_t rows, cols; const void* identity; void* out; cudaStream_t stream;
    template<class T, class Op>
    int operator()(T, Op o) const {
        return launch_reduce_last_dim<T, Op>((const T*)in, rows, cols, load_scalar<T>(identity),
                                             (T*)out, stream, o);
    }
};
struct fold_call {
    const void* parts; int nparts; size_t rows; const void* identity; void* out;
    cudaStream_t stream;
    template<class T, class Op>
    int operator()(T, Op o) const {
        unsigned grid = unsigned((rows + 255) / 256);
        if (grid > 1184) grid = 1184;
        fold_rows_contig_kernel<T, Op><<<grid, 256, 0, stream>>>(
            (const T*)parts, nparts, rows, load_scalar<T>(identity), (T*)out, o);
        SCLX_LAUNCHED();
        return 0;
    }
};

struct alignas(16) bytes16 {
    unsigned long long a, b;
};

}  // namespace

// =====================================================================================
extern "C" {

// internal hooks for the other translation units of this library
unsigned long long sclx_b200_note_launches(unsigned long long n) {
    return g_launches.fetch_add(n, std::memory_order_relaxed) + n;
}
int sclx_b200_set_error(int code, const char* msg) { return fail(code, msg ? msg : ""); }

int sclx_b200_abi_version(void) { return SCLX_B200_ABI_VERSION; }
const char* sclx_b200_last_error(void) { return g_last_error.c_str(); }
unsigned long long sclx_b200_launch_count(void) { return g_launches.load(); }
size_t sclx_b200_dtype_size(int dtype) { return dtype_size(dtype); }

int sclx_b200_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaGetDeviceCount");
        return 0;
    }
    return n;
}

int sclx_b200_set_device(int device) {
    SCLX_CUDA(cudaSetDevice(device));
    return 0;
}

int sclx_b200_partition(size_t last_dim, int ndev, size_t* start, size_t* len) {
    if (ndev < 1 || start == nullptr || len == nullptr)
        return fail(SCLX_E_INVALID_ARGUMENT, "partition: ndev must be >= 1");
    if (last_dim == 0) {
        start[0] = 0;
        len[0]   = 0;
        return 1;
    }
    const size_t split = (last_dim + size_t(ndev) - 1) / size_t(ndev);
    int used           = 0;
    for (size_t s = 0; s < last_dim && used < ndev; s += split, ++used) {
        start[used] = s;
        len[used]   = (last_dim - s < split) ? last_dim - s : split;
    }
    return used;
}

int sclx_b200_malloc(void** ptr, size_t bytes) {
    if (!ptr) return fail(SCLX_E_INVALID_ARGUMENT, "malloc: null out pointer");
    *ptr = nullptr;
    if (bytes == 0) return 0;
    SCLX_CUDA(cudaMalloc(ptr, bytes));
    SCLX_CUDA(cudaMemset(*ptr, 0, bytes));
    return 0;
}
int sclx_b200_free(void* ptr) {
    if (ptr) SCLX_CUDA(cudaFree(ptr));
    return 0;
}
int sclx_b200_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return fail(SCLX_E_INVALID_ARGUMENT, "host_alloc: null out pointer");
    SCLX_CUDA(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocPortable));
    return 0;
}
int sclx_b200_host_free(void* ptr) {
    if (ptr) SCLX_CUDA(cudaFreeHost(ptr));
    return 0;
}
int sclx_b200_memcpy_h2d(void* dst, const void* src, size_t bytes, void* stream) {
    SCLX_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
}
int sclx_b200_memcpy_d2h(void* dst, const void* src, size_t bytes, void* stream) {
    SCLX_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return 0;
}
int sclx_b200_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream) {
    SCLX_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}
int sclx_b200_stream_create(void** stream) {
    cudaStream_t s;
    SCLX_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = s;
    return 0;
}
int sclx_b200_stream_destroy(void* stream) {
    SCLX_CUDA(cudaStreamDestroy((cudaStream_t)stream));
    return 0;
}
int sclx_b200_stream_synchronize(void* stream) {
    SCLX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
int sclx_b200_ipc_get_handle(void* ptr, void* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaIpcMemHandle_t h;
    SCLX_CUDA(cudaIpcGetMemHandle(&h, ptr));
    std::memcpy(handle64, &h, 64);
    return 0;
}
int sclx_b200_ipc_open_handle(const void* handle64, void** ptr) {
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    SCLX_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int sclx_b200_ipc_close(void* ptr) {
    SCLX_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}

int sclx_b200_fill(void* out, size_t n, int elem_bytes, const void* value, void* stream) {
    if (!value || (!out && n)) return fail(SCLX_E_INVALID_ARGUMENT, "fill: null pointer");
    auto s = (cudaStream_t)stream;
    switch (elem_bytes) {
        case 1: return launch_fill((uint8_t*)out, n, load_scalar<uint8_t>(value), s);
        case 2: return launch_fill((uint16_t*)out, n, load_scalar<uint16_t>(value), s);
        case 4: return launch_fill((uint32_t*)out, n, load_scalar<uint32_t>(value), s);
        case 8:
            return launch_fill((unsigned long long*)out, n,
                               load_scalar<unsigned long long>(value), s);
        case 16: return launch_fill((bytes16*)out, n, load_scalar<bytes16>(value), s);
        default: return fail(SCLX_E_UNSUPPORTED, "fill: elem_bytes must be 1,2,4,8 or 16");
    }
}

int sclx_b200_iota(void* out, size_t n, int dtype, const void* start, size_t first_index,
                   void* stream) {
    if (!start || (!out && n)) return fail(SCLX_E_INVALID_ARGUMENT, "iota: null pointer");
    auto s = (cudaStream_t)stream;
    switch (dtype) {
        case SCLX_I32:  // wraps like the unsigned type
        case SCLX_U32:
            return launch_iota((uint32_t*)out, n, load_scalar<uint32_t>(start), first_index, s);
        case SCLX_I64:
        case SCLX_U64:
            return launch_iota((unsigned long long*)out, n,
                               load_scalar<unsigned long long>(start), first_index, s);
        case SCLX_F32:
            return launch_iota((float*)out, n, load_scalar<float>(start), first_index, s);
        case SCLX_F64:
            return launch_iota((double*)out, n, load_scalar<double>(start), first_index, s);
        default: return fail(SCLX_E_UNSUPPORTED, "iota: unknown dtype");
    }
}

int sclx_b200_inclusive_scan(const void* in, void* out, size_t n, int dtype, int op,
                             const void* lower_totals, int n_lower, void* total_out,
                             void* stream) {
    if ((!in || !out) && n)
        return fail(SCLX_E_INVALID_ARGUMENT, "inclusive_scan: null pointer");
    if (n_lower < 0 || (n_lower > 0 && !lower_totals))
        return fail(SCLX_E_INVALID_ARGUMENT, "inclusive_scan: bad carry arguments");
    return dispatch(dtype, op, scan_call{in, out, n, lower_totals, n_lower, total_out, (cudaStream_t)stream});
}

int sclx_b200_reduce(const void* in, size_t n, int dtype, int op, const void* identity,
                     void* out, void* stream) {
    if (!identity || !out || (!in && n))
        return fail(SCLX_E_INVALID_ARGUMENT, "reduce: null pointer");
    return dispatch(dtype, op, reduce_call{in, n, identity, out, (cudaStream_t)stream});
}

int sclx_b200_reduce_last_dim(const void* in, size_t rows, size_t cols, int dtype, int op,
                              const void* identity, void* out, void* stream) {
    if (!identity || (!out && rows) || (!in && rows * cols))
        return fail(SCLX_E_INVALID_ARGUMENT, "reduce_last_dim: null pointer");
    return dispatch(dtype, op, rld_call{in, rows, cols, identity, out, (cudaStream_t)stream});
}

int sclx_b200_fold_rows(const void* parts, int nparts, size_t rows, int dtype, int op,
                        const void* identity, void* out, void* stream) {
    if (nparts < 0 || !identity || (!out && rows) || (!parts && nparts && rows))
        return fail(SCLX_E_INVALID_ARGUMENT, "fold_rows: bad arguments");
    if (rows == 0) return 0;
    return dispatch(dtype, op, fold_call{parts, nparts, rows, identity, out, (cudaStream_t)stream});
}

int sclx_b200_gather(const void* idx, size_t n, int idx_bytes, const void* const* src_parts,
                     int nparts, size_t elems_per_part, int elem_bytes, void* out,
                     void* stream) {
    if (nparts < 1 || nparts > 8 || !src_parts || (n && (!idx || !out)))
        return fail(SCLX_E_INVALID_ARGUMENT, "gather: bad arguments");
    if (nparts > 1 && elems_per_part == 0)
        return fail(SCLX_E_INVALID_ARGUMENT, "gather: elems_per_part must be > 0");
    gather_parts_t parts{};
    for (int p = 0; p < nparts; ++p) parts.part[p] = src_parts[p];
    parts.elems_per_part = elems_per_part ? elems_per_part : ~0ull;
    parts.nparts         = nparts;
    auto s               = (cudaStream_t)stream;
    if (idx_bytes == 8 && elem_bytes == 4)
        return launch_gather<uint32_t, unsigned long long>((const unsigned long long*)idx, n,
                                                           parts, (uint32_t*)out, s);
    if (idx_bytes == 8 && elem_bytes == 8)
        return launch_gather<unsigned long long, unsigned long long>(
            (const unsigned long long*)idx, n, parts, (unsigned long long*)out, s);
    if (idx_bytes == 4 && elem_bytes == 4)
        return launch_gather<uint32_t, uint32_t>((const uint32_t*)idx, n, parts,
                                                 (uint32_t*)out, s);
    if (idx_bytes == 4 && elem_bytes == 8)
        return launch_gather<unsigned long long, uint32_t>((const uint32_t*)idx, n, parts,
                                                           (unsigned long long*)out, s);
    return fail(SCLX_E_UNSUPPORTED, "gather: idx_bytes and elem_bytes must be 4 or 8");
}

// ---- host-buffer entry points ----------------------------------------------------------
// H2D copies are chunked and double-buffered on a copy stream so the kernels run under
// the PCIe transfer; the whole call is bounded by the host link, not by HBM.
namespace {
struct host_pipe {
    cudaStream_t copy = nullptr, exec = nullptr;
    cudaEvent_t filled[2]{}, drained[2]{};
    void* buf[2]{};
    size_t buf_bytes = 0;
    void* small      = nullptr;  // device scratch: partial rows / totals
    size_t small_bytes = 0;
    std::mutex mu;
};
host_pipe g_pipe[kMaxDevices];

int pipe_for_current(host_pipe** out, size_t chunk_bytes, size_t small_bytes) {
    int dev = 0;
    SCLX_CUDA(cudaGetDevice(&dev));
    host_pipe& p = g_pipe[dev];
    if (!p.copy) {
        SCLX_CUDA(cudaStreamCreateWithFlags(&p.copy, cudaStreamNonBlocking));
        SCLX_CUDA(cudaStreamCreateWithFlags(&p.exec, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            SCLX_CUDA(cudaEventCreateWithFlags(&p.filled[i], cudaEventDisableTiming));
            SCLX_CUDA(cudaEventCreateWithFlags(&p.drained[i], cudaEventDisableTiming));
        }
    }
    if (p.buf_bytes < chunk_bytes) {
        SCLX_CUDA(cudaDeviceSynchronize());
        for (int i = 0; i < 2; ++i) {
            if (p.buf[i]) SCLX_CUDA(cudaFree(p.buf[i]));
            SCLX_CUDA(cudaMalloc(&p.buf[i], chunk_bytes));
        }
        p.buf_bytes = chunk_bytes;
    }
    if (p.small_bytes < small_bytes) {
        SCLX_CUDA(cudaDeviceSynchronize());
        if (p.small) SCLX_CUDA(cudaFree(p.small));
        SCLX_CUDA(cudaMalloc(&p.small, small_bytes));
        p.small_bytes = small_bytes;
    }
    *out = &p;
    return 0;
}
constexpr size_t kChunkBytes = size_t(64) << 20;
}  // namespace

int sclx_b200_reduce_last_dim_host(const void* host_in, size_t rows, size_t cols, int dtype,
                                   int op, const void* identity, void* host_out) {
    const size_t es = dtype_size(dtype);
    if (es == 0) return fail(SCLX_E_UNSUPPORTED, "unknown dtype");
    if (!identity || (!host_out && rows) || (!host_in && rows * cols))
        return fail(SCLX_E_INVALID_ARGUMENT, "reduce_last_dim_host: null pointer");
    if (rows == 0) return 0;
    const size_t col_bytes = rows * es;
    size_t cols_per_chunk  = kChunkBytes / col_bytes;
    if (cols_per_chunk == 0) cols_per_chunk = 1;
    const size_t nchunks = cols == 0 ? 0 : (cols + cols_per_chunk - 1) / cols_per_chunk;
    host_pipe* p         = nullptr;
    if (int rc = pipe_for_current(&p, cols_per_chunk * col_bytes, (nchunks + 1) * col_bytes))
        return rc;
    std::lock_guard<std::mutex> lock(p->mu);
    char* part_rows = static_cast<char*>(p->small);           // [nchunks][rows]
    char* final_row = part_rows + nchunks * col_bytes;        // [rows]
    const char* src = static_cast<const char*>(host_in);
    for (size_t c = 0; c < nchunks; ++c) {
        const int b       = int(c & 1);
        const size_t c0   = c * cols_per_chunk;
        const size_t ncol = (cols - c0 < cols_per_chunk) ? cols - c0 : cols_per_chunk;
        if (c >= 2) SCLX_CUDA(cudaStreamWaitEvent(p->copy, p->drained[b], 0));
        SCLX_CUDA(cudaMemcpyAsync(p->buf[b], src + c0 * col_bytes, ncol * col_bytes,
                                  cudaMemcpyHostToDevice, p->copy));
        SCLX_CUDA(cudaEventRecord(p->filled[b], p->copy));
        SCLX_CUDA(cudaStreamWaitEvent(p->exec, p->filled[b], 0));
        if (int rc = sclx_b200_reduce_last_dim(p->buf[b], rows, ncol, dtype, op, identity,
                                               part_rows + c * col_bytes, p->exec))
            return rc;
        SCLX_CUDA(cudaEventRecord(p->drained[b], p->exec));
    }
    if (int rc = sclx_b200_fold_rows(part_rows, int(nchunks), rows, dtype, op, identity,
                                     final_row, p->exec))
        return rc;
    SCLX_CUDA(cudaMemcpyAsync(host_out, final_row, col_bytes, cudaMemcpyDeviceToHost, p->exec));
    SCLX_CUDA(cudaStreamSynchronize(p->exec));
    return 0;
}

int sclx_b200_inclusive_scan_host(const void* host_in, void* host_out, size_t n, int dtype,
                                  int op) {
    const size_t es = dtype_size(dtype);
    if (es == 0) return fail(SCLX_E_UNSUPPORTED, "unknown dtype");
    if ((!host_in || !host_out) && n)
        return fail(SCLX_E_INVALID_ARGUMENT, "inclusive_scan_host: null pointer");
    if (n == 0) return 0;
    const size_t per_chunk = kChunkBytes / es;
    const size_t nchunks   = (n + per_chunk - 1) / per_chunk;
    host_pipe* p           = nullptr;
    if (int rc = pipe_for_current(&p, per_chunk * es, 2 * 16)) return rc;
    std::lock_guard<std::mutex> lock(p->mu);
    const char* src = static_cast<const char*>(host_in);
    char* dst       = static_cast<char*>(host_out);
    char* totals    = static_cast<char*>(p->small);  // ping-pong carried totals
    for (size_t c = 0; c < nchunks; ++c) {
        const int b     = int(c & 1);
        const size_t e0 = c * per_chunk;
        const size_t ne = (n - e0 < per_chunk) ? n - e0 : per_chunk;
        if (c >= 2) SCLX_CUDA(cudaStreamWaitEvent(p->copy, p->drained[b], 0));
        SCLX_CUDA(cudaMemcpyAsync(p->buf[b], src + e0 * es, ne * es, cudaMemcpyHostToDevice,
                                  p->copy));
        SCLX_CUDA(cudaEventRecord(p->filled[b], p->copy));
        SCLX_CUDA(cudaStreamWaitEvent(p->exec, p->filled[b], 0));
        // chunk c is carried by the (already carried) last value of chunk c-1
        if (int rc = sclx_b200_inclusive_scan(
                p->buf[b], p->buf[b], ne, dtype, op, c ? totals + ((c - 1) & 1) * 16 : nullptr,
                c ? 1 : 0, totals + (c & 1) * 16, p->exec))
            return rc;
        SCLX_CUDA(cudaMemcpyAsync(dst + e0 * es, p->buf[b], ne * es, cudaMemcpyDeviceToHost,
                                  p->exec));
        SCLX_CUDA(cudaEventRecord(p->drained[b], p->exec));
    }
    SCLX_CUDA(cudaStreamSynchronize(p->exec));
    return 0;
}

}  // extern "C"
