// scalix-b200 runtime: the part of the hot path that is NOT a kernel.
//
//  (1) memory manager — replaces the cudaMallocManaged + cudaMemAdvise/prefetch scheme of
//      /root/reference scalix/array.cuh:241-248,517-615.  Every array is ONE contiguous
//      virtual range (CUDA VMM) whose pages are explicit per-device physical allocations
//      placed according to the last-dimension split and mapped read/write into every
//      device, so a single pointer means the same thing on all GPUs: local accesses hit
//      HBM, remote ones travel over NVLink/NVSwitch as plain loads/stores, nothing faults.
//      Host access (the reference's examples dereference array memory on the CPU,
//      examples/inclusive_scan.cu:87-92) is served by a software mirror overlaid on the
//      SAME virtual addresses: host pages start PROT_NONE, are filled on first touch from
//      the device, written back before the next device access.  No page migration, no
//      driver heuristics.
//  (2) launch plumbing — replaces stream-per-launch + std::async-per-kernel
//      (/root/reference scalix/detail/execute_kernel.cuh:110-115,197-199,
//      scalix/execute_kernel.cuh:506-513): one persistent non-blocking stream per logical
//      device, launch ordering through events, explicit synchronisation only where the
//      API promises it.
//  (3) logical-device emulation — the equivalent of SCALIX_EMULATE_MULTIDEVICE
//      (/root/reference CMakeLists.txt:31-44): N logical devices on one physical GPU.
#include "scalix_b200.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <csignal>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <ucontext.h>
#include <unistd.h>

extern "C" int sclx_b200_set_error(int code, const char* msg);

namespace {

// ---- driver entry points (resolved at run time: the library must load without libcuda) ----
struct driver_api {
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
    CUresult (*StreamWriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int) = nullptr;  // optional
    bool ok = false;
};
driver_api g_drv;

template<class F>
bool resolve(const char* name, F& fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || !p) return false;
    fn = reinterpret_cast<F>(p);
    return true;
}

int fail(int code, const std::string& msg) { return sclx_b200_set_error(code, msg.c_str()); }
int cu_fail(CUresult r, const char* what) {
    const char* s = nullptr;
    if (g_drv.GetErrorString) g_drv.GetErrorString(r, &s);
    return fail(SCLX_E_RUNTIME, std::string(what) + ": " + (s ? s : "CUDA driver error"));
}
int rt_fail(cudaError_t e, const char* what) {
    return fail(int(e), std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define RT(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return rt_fail(e__, #call); } while (0)
#define CU(call) do { CUresult r__ = (call); if (r__ != CUDA_SUCCESS) return cu_fail(r__, #call); } while (0)

// ---- global state --------------------------------------------------------------------------
constexpr size_t kMirrorChunk = size_t(64) << 10;  // host-mirror coherence granule
constexpr size_t kPinThreshold = size_t(64) << 20;  // bulk CPU prefetches from this size on pin the mirror

enum chunk_state : unsigned char { kInvalid = 0, kClean = 1, kDirty = 2 };

struct allocation {
    char* base      = nullptr;  // device VA == host overlay VA
    size_t bytes    = 0;        // user-visible size
    size_t va_bytes = 0;        // reserved size (granularity multiple)
    struct run {                // one physical allocation
        size_t offset, bytes;
        int logical_device;
        CUmemGenericAllocationHandle handle;
    };
    std::vector<run> runs;
    // host mirror
    int memfd       = -1;
    char* alias     = nullptr;  // always-RW view of the mirror pages (runtime only)
    bool overlay_ok = false;
    std::vector<unsigned char> state;  // per mirror chunk
    size_t n_valid = 0;
    // a bulk prefetch to the CPU pins the mirror pages (cudaHostRegister) so the copies in both
    // directions run at the PCIe rate; pinned pages are kept until the array is freed
    bool pinned = false;
    // sequential host access grows the span a fault brings in: 64 KiB, 128 KiB, ... 16 MiB
    size_t fault_next = ~size_t(0), fault_span = 1;
};

// workspaces of the framework-owned kernels, leased per call (never shared by two calls in
// flight): the reference keeps its partials and mutexes function-local
// (/root/reference scalix/algorithm/reduce_last_dim.cuh:180-211, inclusive_scan.cuh:75-123)
struct kernel_workspace {
    int logical        = 0;
    void* partials     = nullptr; size_t partial_bytes = 0;
    unsigned* ticket   = nullptr;  // 256 zeroed bytes: [0] reduce ticket, [32..] scan tile tickets
    // look-back descriptors, one buffer and one epoch PER WORD SIZE (8-byte words for elements
    // of <= 4 bytes, 16-byte words for 8-byte elements): a stale word of the other layout must
    // never be read as a valid {epoch|state, value} pair
    void* desc[2]      = {nullptr, nullptr};
    size_t desc_bytes[2] = {0, 0};
    unsigned epoch[2]  = {0, 0};
};

struct logical_device {
    int physical        = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t done    = nullptr;  // recorded (lazily) after the latest launch group
    // launch-group sequence numbers (guarded by g_rt.mu): the latest group that queued work on
    // this device, and the latest one `done` covers.  `done` is only recorded when a later
    // group has to wait for it, never in the blocking-call pattern where the host syncs first.
    unsigned long long last_launch_seq = 0, recorded_seq = 0;
    unsigned long long host_synced_seq = 0;  // groups <= this were waited for by a host sync of this device
    cudaEvent_t gate    = nullptr;  // recorded by the hub of a launch group once every device's `done` has fired
    int sm_count        = 0;
    std::vector<kernel_workspace*> free_ws;  // leases not in use (guarded by g_rt.mu)
    void* exchange = nullptr; size_t exchange_bytes = 0;  // peer-visible gather-exchange buffers
};

struct runtime {
    std::recursive_mutex mu;
    bool initialised = false;
    int physical_count = 0;
    int emulated       = 0;  // 0 = use physical devices
    std::vector<logical_device> dev;
    std::map<char*, allocation*> allocs;  // keyed by base
    size_t granularity = size_t(2) << 20;
    std::atomic<size_t> host_valid_chunks{0};
    // launch groups issued / covered by a full host synchronisation (monotonic): work is
    // outstanding while launch_seq != synced_seq.  Two counters instead of a flag so that a
    // sync on one host thread cannot hide a launch another thread issued meanwhile.
    std::atomic<unsigned long long> launch_seq{0}, synced_seq{0};
    bool handler_installed = false;
    struct sigaction old_action {};
    struct scratch_block { void* ptr; size_t bytes; bool busy; };
    std::vector<scratch_block> scratch;  // small peer-visible blocks, leased per call
    unsigned long long faults = 0, flushed_chunks = 0;
    // freed arrays kept mapped for reuse (VA + physical pages + peer access + mirror intact):
    // creating a multi-GiB peer-mapped range costs tens of ms on a multi-GPU node, reusing one
    // costs its zero-fill.  Oldest first; bounded by pool_cap bytes (SCLX_B200_POOL_BYTES).
    std::vector<allocation*> pool;
    size_t pool_bytes = 0, pool_cap = size_t(32) << 30;
    unsigned long long pool_hits = 0, pool_misses = 0;
};
runtime g_rt;

// Completion words: after the kernels of a launch group the issuing thread queues a 32-bit
// stream write of the device's next ticket into a word of pinned, mapped host memory; waiting
// for a device = polling its word (one cache line of host memory) instead of a
// cudaStreamSynchronize per device, which on a multi-GPU node costs 3-5 us EACH, one after the
// other, on the calling thread (8 B200: 40 of the 655 us of a reduce_last_dim call).
constexpr int kMaxSignalDevices = 64;
struct device_signal {
    std::mutex mu;                 // ticket order == stream order
    unsigned issued = 0;           // last ticket queued (guarded by mu; read racily by waiters)
    volatile unsigned* word = nullptr;
};
device_signal g_signal[kMaxSignalDevices];

int ensure_init() {
    if (g_rt.initialised) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    if (g_rt.initialised) return 0;
    int n = 0;
    RT(cudaGetDeviceCount(&n));
    if (n < 1) return fail(SCLX_E_RUNTIME, "no CUDA device");
    g_rt.physical_count = n;
    bool ok = resolve("cuMemGetAllocationGranularity", g_drv.MemGetAllocationGranularity)
           && resolve("cuMemAddressReserve", g_drv.MemAddressReserve)
           && resolve("cuMemAddressFree", g_drv.MemAddressFree)
           && resolve("cuMemCreate", g_drv.MemCreate) && resolve("cuMemRelease", g_drv.MemRelease)
           && resolve("cuMemMap", g_drv.MemMap) && resolve("cuMemUnmap", g_drv.MemUnmap)
           && resolve("cuMemSetAccess", g_drv.MemSetAccess)
           && resolve("cuGetErrorString", g_drv.GetErrorString);
    if (!ok) return fail(SCLX_E_RUNTIME, "CUDA virtual memory management entry points unavailable");
    g_drv.ok = true;
    if (!getenv("SCLX_B200_NO_STREAM_SIGNALS")) resolve("cuStreamWriteValue32", g_drv.StreamWriteValue32);
    if (const char* e = getenv("SCLX_B200_POOL_BYTES")) g_rt.pool_cap = (size_t)strtoull(e, nullptr, 10);
    if (g_rt.emulated == 0) {
        if (const char* e = getenv("SCLX_B200_EMULATE_DEVICES")) {
            int v = atoi(e);
            if (v > 0) g_rt.emulated = v;
        }
    }
    const int logical = g_rt.emulated > 0 ? g_rt.emulated : n;
    int cur = 0;
    cudaGetDevice(&cur);
    g_rt.dev.resize(logical);
    for (int d = 0; d < logical; ++d) {
        logical_device& ld = g_rt.dev[d];
        ld.physical        = g_rt.emulated > 0 ? 0 : d;
        RT(cudaSetDevice(ld.physical));
        RT(cudaFree(nullptr));
        RT(cudaStreamCreateWithFlags(&ld.stream, cudaStreamNonBlocking));
        RT(cudaEventCreateWithFlags(&ld.done, cudaEventDisableTiming));
        RT(cudaEventCreateWithFlags(&ld.gate, cudaEventDisableTiming));
        RT(cudaDeviceGetAttribute(&ld.sm_count, cudaDevAttrMultiProcessorCount, ld.physical));
        if (d < kMaxSignalDevices && g_drv.StreamWriteValue32) {
            void* w = nullptr;
            if (cudaHostAlloc(&w, 64, cudaHostAllocPortable | cudaHostAllocMapped) == cudaSuccess) {
                memset(w, 0, 64);
                g_signal[d].word = static_cast<volatile unsigned*>(w);
            } else cudaGetLastError();
        }
    }
    if (!getenv("SCLX_B200_POOL_BYTES")) {
        // parked arrays hold device memory other allocators in the process cannot see: keep the
        // default cap to an eighth of the device
        size_t free_b = 0, total_b = 0;
        RT(cudaSetDevice(g_rt.dev[0].physical));
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && total_b / 8 < g_rt.pool_cap)
            g_rt.pool_cap = total_b / 8;
    }
    if (g_rt.emulated == 0) {
        // NVLink/NVSwitch peer access for cudaMalloc'd workspaces (VMM ranges carry their own ACLs)
        for (int a = 0; a < n; ++a)
            for (int b = 0; b < n; ++b) {
                if (a == b) continue;
                int can = 0;
                cudaDeviceCanAccessPeer(&can, a, b);
                if (can) {
                    cudaSetDevice(a);
                    cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
                    if (e != cudaSuccess) cudaGetLastError();
                }
            }
    }
    RT(cudaSetDevice(cur));
    CUmemAllocationProp prop{};
    prop.type          = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id   = 0;
    size_t g           = 0;
    CU(g_drv.MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (g < kMirrorChunk) g = size_t(2) << 20;
    g_rt.granularity = g;
    g_rt.initialised = true;
    return 0;
}

allocation* find_alloc(const void* p) {
    char* a = (char*)p;
    auto it = g_rt.allocs.upper_bound(a);
    if (it == g_rt.allocs.begin()) return nullptr;
    --it;
    allocation* al = it->second;
    if (a >= al->base && a < al->base + al->va_bytes) return al;
    return nullptr;
}

int physical_of(int logical) { return g_rt.dev[logical].physical; }

// ---- host mirror ---------------------------------------------------------------------------
bool work_outstanding() { return g_rt.launch_seq.load() != g_rt.synced_seq.load(); }

// every launch group numbered <= seq has been waited for by the host
void note_synced(unsigned long long seq) {
    unsigned long long cur = g_rt.synced_seq.load();
    while (cur < seq && !g_rt.synced_seq.compare_exchange_weak(cur, seq)) {}
}

int sync_all_streams() {
    int cur = 0;
    cudaGetDevice(&cur);
    const unsigned long long seq = g_rt.launch_seq.load();
    for (auto& ld : g_rt.dev) {
        RT(cudaSetDevice(ld.physical));
        RT(cudaStreamSynchronize(ld.stream));
    }
    // work the user issued outside the runtime (thrust on raw pointers, default stream)
    for (int p = 0; p < (g_rt.emulated ? 1 : g_rt.physical_count); ++p) {
        RT(cudaSetDevice(p));
        RT(cudaDeviceSynchronize());
    }
    RT(cudaSetDevice(cur));
    note_synced(seq);
    return 0;
}

int owner_of(const allocation& al, size_t offset) {
    for (auto& r : al.runs)
        if (offset >= r.offset && offset < r.offset + r.bytes) return r.logical_device;
    return al.runs.empty() ? 0 : al.runs.back().logical_device;
}

// a copy that touches device memory of `al` in [off, off+len), split at physical-run
// boundaries (each piece lives on one device)
template<class F>
int for_each_piece(const allocation& al, size_t off, size_t len, F&& f) {
    size_t done = 0;
    while (done < len) {
        size_t o = off + done, piece = len - done;
        for (auto& r : al.runs)
            if (o >= r.offset && o < r.offset + r.bytes) {
                piece = std::min(piece, r.offset + r.bytes - o);
                break;
            }
        if (int rc = f(o, piece)) return rc;
        done += piece;
    }
    return 0;
}

// One transfer between the mirror and device memory, [off, off+len) of `al`, split at
// physical-run boundaries.  A pinned mirror moves each piece asynchronously on its owner's
// stream (several GPUs = several PCIe links at once) and waits for all of them at the end; a
// pageable one uses blocking copies.
int wait_for_copies(std::vector<int>& used) {
    int cur = 0, rc = 0;
    cudaGetDevice(&cur);
    for (int d : used) {
        cudaSetDevice(g_rt.dev[d].physical);
        cudaError_t e = cudaStreamSynchronize(g_rt.dev[d].stream);
        if (e != cudaSuccess && rc == 0) rc = rt_fail(e, "cudaStreamSynchronize");
    }
    cudaSetDevice(cur);
    used.clear();
    return rc;
}

// `pending`: when given, the copies of a pinned mirror are only QUEUED and their devices added
// to it; the caller overlaps host work with the DMA and calls wait_for_copies() afterwards
int copy_pieces(allocation& al, size_t off, size_t len, bool to_host, std::vector<int>* pending = nullptr) {
    if (len == 0) return 0;
    if (!al.pinned)
        return for_each_piece(al, off, len, [&](size_t o, size_t piece) -> int {
            if (to_host) RT(cudaMemcpy(al.alias + o, al.base + o, piece, cudaMemcpyDeviceToHost));
            else RT(cudaMemcpy(al.base + o, al.alias + o, piece, cudaMemcpyHostToDevice));
            return 0;
        });
    int cur = 0;
    cudaGetDevice(&cur);
    std::vector<int> used;
    int rc = for_each_piece(al, off, len, [&](size_t o, size_t piece) -> int {
        const int owner    = owner_of(al, o);
        logical_device& ld = g_rt.dev[owner];
        RT(cudaSetDevice(ld.physical));
        if (to_host) RT(cudaMemcpyAsync(al.alias + o, al.base + o, piece, cudaMemcpyDeviceToHost, ld.stream));
        else RT(cudaMemcpyAsync(al.base + o, al.alias + o, piece, cudaMemcpyHostToDevice, ld.stream));
        if (std::find(used.begin(), used.end(), owner) == used.end()) used.push_back(owner);
        return 0;
    });
    cudaSetDevice(cur);
    if (pending) {
        for (int d : used)
            if (std::find(pending->begin(), pending->end(), d) == pending->end()) pending->push_back(d);
        return rc;
    }
    const int wrc = wait_for_copies(used);
    return rc ? rc : wrc;
}

// bring chunks [c0, c1) of the host mirror up to date (device -> host); `for_write` leaves them
// writable and marked dirty (they go back to the device before its next access)
int fetch_chunks(allocation& al, size_t c0, size_t c1, bool for_write) {
    // coalesce invalid chunks into runs so a bulk prefetch is one copy per physical run
    size_t c = c0;
    bool synced = false;
    while (c < c1) {
        if (al.state[c] != kInvalid) {
            size_t e = c;
            while (e < c1 && al.state[e] != kInvalid) ++e;
            if (for_write) {
                bool any_clean = false;
                for (size_t k = c; k < e; ++k)
                    if (al.state[k] == kClean) { al.state[k] = kDirty; any_clean = true; }
                if (any_clean) mprotect(al.base + c * kMirrorChunk, (e - c) * kMirrorChunk, PROT_READ | PROT_WRITE);
            }
            c = e;
            continue;
        }
        size_t e = c;
        while (e < c1 && al.state[e] == kInvalid) ++e;
        if (!synced) {
            if (int rc = sync_all_streams()) return rc;
            synced = true;
        }
        size_t off = c * kMirrorChunk, len = (e - c) * kMirrorChunk;
        if (off + len > al.va_bytes) len = al.va_bytes - off;
        if (int rc = copy_pieces(al, off, len, /*to_host=*/true)) return rc;
        for (size_t k = c; k < e; ++k) al.state[k] = for_write ? kDirty : kClean;
        mprotect(al.base + off, (e - c) * kMirrorChunk, for_write ? (PROT_READ | PROT_WRITE) : PROT_READ);
        al.n_valid += e - c;
        g_rt.host_valid_chunks.fetch_add(e - c);
        c = e;
    }
    return 0;
}

// write dirty chunks back and drop every host-valid chunk of `al`
int flush_and_invalidate(allocation& al) {
    if (al.n_valid == 0) return 0;
    const size_t nchunks = al.state.size();
    // pinned mirror: the write-back DMA reads physical pages, so the page-table work of the
    // invalidation below (mprotect over gigabytes is tens of milliseconds) runs under it
    std::vector<int> pending;
    size_t c = 0;
    while (c < nchunks) {
        if (al.state[c] == kInvalid) { ++c; continue; }
        size_t e = c;
        while (e < nchunks && al.state[e] != kInvalid) ++e;
        // write back dirty sub-runs
        size_t d = c;
        while (d < e) {
            if (al.state[d] != kDirty) { ++d; continue; }
            size_t f = d;
            while (f < e && al.state[f] == kDirty) ++f;
            size_t off = d * kMirrorChunk, len = (f - d) * kMirrorChunk;
            if (off + len > al.va_bytes) len = al.va_bytes - off;
            if (int rc = copy_pieces(al, off, len, /*to_host=*/false, &pending)) return rc;
            g_rt.flushed_chunks += f - d;
            d = f;
        }
        size_t off = c * kMirrorChunk, len = (e - c) * kMirrorChunk;
        mprotect(al.base + off, len, PROT_NONE);
        // pinned pages stay (the registration holds them); pageable ones go back to the kernel
        if (al.memfd >= 0 && !al.pinned) fallocate(al.memfd, FALLOC_FL_PUNCH_HOLE | FALLOC_FL_KEEP_SIZE, off, len);
        for (size_t k = c; k < e; ++k) al.state[k] = kInvalid;
        g_rt.host_valid_chunks.fetch_sub(e - c);
        c = e;
    }
    al.n_valid    = 0;
    al.fault_next = ~size_t(0);
    al.fault_span = 1;
    return wait_for_copies(pending);
}

constexpr size_t kMaxFaultSpan = 256;  // chunks: 16 MiB

void on_segv(int sig, siginfo_t* si, void* uctx) {
    char* addr = (char*)si->si_addr;
    bool handled = false;
    {
        std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
        allocation* al = find_alloc(addr);
        if (al && al->overlay_ok) {
            bool write = true;
#if defined(__x86_64__)
            write = (((ucontext_t*)uctx)->uc_mcontext.gregs[REG_ERR] & 2) != 0;
#endif
            const size_t nchunks = al->state.size();
            size_t c = size_t(addr - al->base) / kMirrorChunk;
            // a fault right behind the previous one is a sequential sweep: double the span
            size_t span = 1;
            if (c == al->fault_next) span = std::min(al->fault_span * 2, kMaxFaultSpan);
            const size_t c1 = std::min(c + span, nchunks);
            ++g_rt.faults;
            if (fetch_chunks(*al, c, c1, write) == 0) {
                handled        = true;
                al->fault_next = c1;
                al->fault_span = span;
            }
        }
    }
    if (handled) return;
    // not ours: hand over to whoever was installed before (or die the normal way)
    if (g_rt.old_action.sa_flags & SA_SIGINFO) {
        if (g_rt.old_action.sa_sigaction) { g_rt.old_action.sa_sigaction(sig, si, uctx); return; }
    } else if (g_rt.old_action.sa_handler != SIG_DFL && g_rt.old_action.sa_handler != SIG_IGN) {
        g_rt.old_action.sa_handler(sig);
        return;
    }
    signal(SIGSEGV, SIG_DFL);
}

void install_handler() {
    if (g_rt.handler_installed) return;
    struct sigaction sa {};
    sa.sa_sigaction = on_segv;
    sa.sa_flags     = SA_SIGINFO | SA_NODEFER;
    sigemptyset(&sa.sa_mask);
    sigaction(SIGSEGV, &sa, &g_rt.old_action);
    g_rt.handler_installed = true;
}

void setup_overlay(allocation& al) {
    if (getenv("SCLX_B200_NO_HOST_MIRROR")) return;
    al.memfd = memfd_create("sclx_b200_mirror", MFD_CLOEXEC);
    if (al.memfd < 0) return;
    if (ftruncate(al.memfd, (off_t)al.va_bytes) != 0) { close(al.memfd); al.memfd = -1; return; }
    void* alias = mmap(nullptr, al.va_bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_NORESERVE, al.memfd, 0);
    if (alias == MAP_FAILED) { close(al.memfd); al.memfd = -1; return; }
    void* over = mmap(al.base, al.va_bytes, PROT_NONE, MAP_SHARED | MAP_FIXED | MAP_NORESERVE, al.memfd, 0);
    if (over != (void*)al.base) {
        munmap(alias, al.va_bytes);
        close(al.memfd);
        al.memfd = -1;
        return;
    }
    al.alias      = (char*)alias;
    al.overlay_ok = true;
    al.state.assign((al.va_bytes + kMirrorChunk - 1) / kMirrorChunk, kInvalid);
    install_handler();
}

// ---- physical placement ----------------------------------------------------------------------
// page k belongs to the partition that contains byte k*granularity
int build_runs(allocation& al, int nparts, const int* devices, const size_t* starts) {
    const size_t g = g_rt.granularity;
    auto page_ceil = [&](size_t b) { return (b + g - 1) / g; };
    const size_t total_pages = al.va_bytes / g;
    size_t first = 0;
    for (int p = 0; p < nparts; ++p) {
        size_t next = (p + 1 < nparts) ? std::min(page_ceil(starts[p + 1]), total_pages) : total_pages;
        if (next > first) {
            allocation::run r{first * g, (next - first) * g, devices[p], 0};
            al.runs.push_back(r);
        }
        first = std::max(first, next);
    }
    return 0;
}

int map_runs(allocation& al, bool zero) {
    int cur = 0;
    cudaGetDevice(&cur);
    std::vector<CUmemAccessDesc> acc;
    const int nphys = g_rt.emulated ? 1 : g_rt.physical_count;
    for (int p = 0; p < nphys; ++p) {
        CUmemAccessDesc d{};
        d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        d.location.id   = p;
        d.flags         = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        acc.push_back(d);
    }
    for (auto& r : al.runs) {
        CUmemAllocationProp prop{};
        prop.type          = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id   = physical_of(r.logical_device);
        CU(g_drv.MemCreate(&r.handle, r.bytes, &prop, 0));
        CU(g_drv.MemMap((CUdeviceptr)(al.base + r.offset), r.bytes, 0, r.handle, 0));
    }
    // every device may read and write every partition (peer loads/stores over NVLink)
    CUresult ar = g_drv.MemSetAccess((CUdeviceptr)al.base, al.va_bytes, acc.data(), acc.size());
    if (ar != CUDA_SUCCESS) {
        // no peer path between some pair: fall back to owner-only access per run
        for (auto& r : al.runs) {
            CUmemAccessDesc d{};
            d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            d.location.id   = physical_of(r.logical_device);
            d.flags         = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
            CU(g_drv.MemSetAccess((CUdeviceptr)(al.base + r.offset), r.bytes, &d, 1));
        }
    }
    if (zero) {
        for (auto& r : al.runs) {
            logical_device& ld = g_rt.dev[r.logical_device];
            RT(cudaSetDevice(ld.physical));
            RT(cudaMemsetAsync(al.base + r.offset, 0, r.bytes, ld.stream));
        }
        // the caller may touch the memory from ANY stream (thrust on raw pointers, cudaMemcpy)
        // right after creation: the zero-fill must not still be in flight
        for (auto& r : al.runs) {
            logical_device& ld = g_rt.dev[r.logical_device];
            RT(cudaSetDevice(ld.physical));
            RT(cudaStreamSynchronize(ld.stream));
        }
    }
    RT(cudaSetDevice(cur));
    return 0;
}

int unmap_runs(allocation& al) {
    for (auto& r : al.runs) {
        CU(g_drv.MemUnmap((CUdeviceptr)(al.base + r.offset), r.bytes));
        CU(g_drv.MemRelease(r.handle));
    }
    al.runs.clear();
    return 0;
}

// tear an allocation down completely (mappings, physical pages, mirror, reservation)
void unpin_mirror(allocation& al) {
    if (!al.pinned) return;
    if (cudaHostUnregister(al.alias) != cudaSuccess) cudaGetLastError();
    al.pinned = false;
}

int destroy_allocation(allocation* al) {
    unpin_mirror(*al);
    if (al->overlay_ok) {
        // give the range back to a plain reservation before the driver unmaps it
        mmap(al->base, al->va_bytes, PROT_NONE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_FIXED | MAP_NORESERVE, -1, 0);
        munmap(al->alias, al->va_bytes);
        close(al->memfd);
    }
    int rc = unmap_runs(*al);
    if (!rc) {
        CUresult r = g_drv.MemAddressFree((CUdeviceptr)al->base, al->va_bytes);
        if (r != CUDA_SUCCESS) rc = cu_fail(r, "cuMemAddressFree");
    }
    delete al;
    return rc;
}

// forget every host-valid chunk WITHOUT writing anything back (the array is dead)
void drop_mirror(allocation& al) {
    if (!al.overlay_ok) return;
    const bool was_pinned = al.pinned;
    unpin_mirror(al);
    if (was_pinned && al.memfd >= 0) fallocate(al.memfd, FALLOC_FL_PUNCH_HOLE | FALLOC_FL_KEEP_SIZE, 0, (off_t)al.va_bytes);
    al.fault_next = ~size_t(0);
    al.fault_span = 1;
    const size_t nchunks = al.state.size();
    size_t c = 0;
    while (c < nchunks) {
        if (al.state[c] == kInvalid) { ++c; continue; }
        size_t e = c;
        while (e < nchunks && al.state[e] != kInvalid) ++e;
        const size_t off = c * kMirrorChunk, len = (e - c) * kMirrorChunk;
        mprotect(al.base + off, len, PROT_NONE);
        if (al.memfd >= 0) fallocate(al.memfd, FALLOC_FL_PUNCH_HOLE | FALLOC_FL_KEEP_SIZE, off, len);
        for (size_t k = c; k < e; ++k) al.state[k] = kInvalid;
        c = e;
    }
    al.n_valid = 0;
}

int zero_runs(allocation& al) {
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto& r : al.runs) {
        logical_device& ld = g_rt.dev[r.logical_device];
        RT(cudaSetDevice(ld.physical));
        RT(cudaMemsetAsync(al.base + r.offset, 0, r.bytes, ld.stream));
    }
    for (auto& r : al.runs) {
        logical_device& ld = g_rt.dev[r.logical_device];
        RT(cudaSetDevice(ld.physical));
        RT(cudaStreamSynchronize(ld.stream));
    }
    RT(cudaSetDevice(cur));
    return 0;
}

void pool_purge(size_t keep_bytes) {
    while (!g_rt.pool.empty() && g_rt.pool_bytes > keep_bytes) {
        allocation* al = g_rt.pool.front();
        g_rt.pool.erase(g_rt.pool.begin());
        g_rt.pool_bytes -= al->va_bytes;
        destroy_allocation(al);
    }
}

// a pooled allocation with exactly this placement, or nullptr
allocation* pool_take(size_t va_bytes, const std::vector<allocation::run>& want) {
    for (size_t i = g_rt.pool.size(); i-- > 0;) {
        allocation* al = g_rt.pool[i];
        if (al->va_bytes != va_bytes || al->runs.size() != want.size()) continue;
        bool same = true;
        for (size_t k = 0; same && k < want.size(); ++k)
            same = al->runs[k].offset == want[k].offset && al->runs[k].bytes == want[k].bytes
                && physical_of(al->runs[k].logical_device) == physical_of(want[k].logical_device);
        if (!same) continue;
        for (size_t k = 0; k < want.size(); ++k) al->runs[k].logical_device = want[k].logical_device;
        g_rt.pool.erase(g_rt.pool.begin() + (long)i);
        g_rt.pool_bytes -= al->va_bytes;
        return al;
    }
    return nullptr;
}

int validate_parts(size_t bytes, int nparts, const int* devices, const size_t* starts) {
    if (nparts < 1 || !devices || !starts) return fail(SCLX_E_INVALID_ARGUMENT, "array: bad partition list");
    if (starts[0] != 0) return fail(SCLX_E_INVALID_ARGUMENT, "array: first partition must start at byte 0");
    for (int p = 0; p < nparts; ++p) {
        if (devices[p] < 0 || devices[p] >= (int)g_rt.dev.size())
            return fail(SCLX_E_INVALID_ARGUMENT, "array: device id out of range");
        if (p && starts[p] < starts[p - 1]) return fail(SCLX_E_INVALID_ARGUMENT, "array: partition starts must ascend");
        if (starts[p] > bytes) return fail(SCLX_E_INVALID_ARGUMENT, "array: partition start beyond the array");
    }
    return 0;
}

}  // namespace

// =================================================================================================
extern "C" {

int sclx_b200_rt_init(void) { return ensure_init(); }

int sclx_b200_rt_set_emulated_devices(int n) {
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    if (g_rt.initialised) {
        if ((g_rt.emulated > 0 ? g_rt.emulated : 0) == (n > 0 ? n : 0)) return 0;
        return fail(SCLX_E_RUNTIME, "emulated device count must be set before the runtime starts");
    }
    g_rt.emulated = n > 0 ? n : 0;
    return 0;
}

int sclx_b200_rt_device_count(void) {
    if (ensure_init()) return 0;
    return (int)g_rt.dev.size();
}

int sclx_b200_rt_physical_device(int logical) {
    if (ensure_init() || logical < 0 || logical >= (int)g_rt.dev.size()) return -1;
    return g_rt.dev[logical].physical;
}

void* sclx_b200_rt_stream(int logical) {
    if (ensure_init() || logical < 0 || logical >= (int)g_rt.dev.size()) return nullptr;
    return g_rt.dev[logical].stream;
}

int sclx_b200_rt_sm_count(int logical) {
    if (ensure_init() || logical < 0 || logical >= (int)g_rt.dev.size()) return 0;
    return g_rt.dev[logical].sm_count;
}

int sclx_b200_array_create(size_t bytes, int nparts, const int* devices, const size_t* byte_starts,
                           void** base) {
    if (!base) return fail(SCLX_E_INVALID_ARGUMENT, "array_create: null out pointer");
    *base = nullptr;
    if (int rc = ensure_init()) return rc;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    if (int rc = validate_parts(bytes, nparts, devices, byte_starts)) return rc;
    const size_t g  = g_rt.granularity;
    allocation* al  = new allocation;
    al->bytes       = bytes;
    al->va_bytes    = std::max<size_t>(g, (bytes + g - 1) / g * g);
    build_runs(*al, nparts, devices, byte_starts);
    if (allocation* pooled = pool_take(al->va_bytes, al->runs)) {
        delete al;
        pooled->bytes = bytes;
        if (int rc = zero_runs(*pooled)) { destroy_allocation(pooled); return rc; }
        ++g_rt.pool_hits;
        g_rt.allocs[pooled->base] = pooled;
        *base                     = pooled->base;
        return 0;
    }
    ++g_rt.pool_misses;
    CUdeviceptr va  = 0;
    CUresult r      = g_drv.MemAddressReserve(&va, al->va_bytes, g, 0, 0);
    if (r != CUDA_SUCCESS) { delete al; return cu_fail(r, "cuMemAddressReserve"); }
    al->base = (char*)va;
    int rc   = map_runs(*al, true);
    if (rc && !g_rt.pool.empty()) {
        // out of device memory with arrays parked in the pool: give them back and retry once
        for (auto& rr : al->runs)
            if (rr.handle) { g_drv.MemUnmap((CUdeviceptr)(al->base + rr.offset), rr.bytes); g_drv.MemRelease(rr.handle); rr.handle = 0; }
        pool_purge(0);
        rc = map_runs(*al, true);
    }
    if (rc) {
        for (auto& rr : al->runs)
            if (rr.handle) { g_drv.MemUnmap((CUdeviceptr)(al->base + rr.offset), rr.bytes); g_drv.MemRelease(rr.handle); }
        al->runs.clear();
        g_drv.MemAddressFree(va, al->va_bytes);
        delete al;
        return rc;
    }
    setup_overlay(*al);
    g_rt.allocs[al->base] = al;
    *base                 = al->base;
    return 0;
}

int sclx_b200_array_free(void* base) {
    if (!base) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    auto it = g_rt.allocs.find((char*)base);
    if (it == g_rt.allocs.end()) return fail(SCLX_E_INVALID_ARGUMENT, "array_free: unknown pointer");
    allocation* al = it->second;
    if (work_outstanding()) {  // kernels in flight may still use it
        const unsigned long long synced = g_rt.synced_seq.load();
        bool busy = false;
        for (auto& ld : g_rt.dev) busy |= ld.last_launch_seq > synced && ld.last_launch_seq > ld.host_synced_seq;
        if (busy) sync_all_streams();
    }
    if (al->n_valid) {
        g_rt.host_valid_chunks.fetch_sub(al->n_valid);
        al->n_valid = 0;
    }
    g_rt.allocs.erase(it);
    if (g_rt.pool_cap > 0 && al->va_bytes <= g_rt.pool_cap && !al->runs.empty()) {
        drop_mirror(*al);
        g_rt.pool.push_back(al);
        g_rt.pool_bytes += al->va_bytes;
        pool_purge(g_rt.pool_cap);
        return 0;
    }
    return destroy_allocation(al);
}

// move the physical pages to a new split, keeping the pointer (and the data)
int sclx_b200_array_repartition(void* base, int nparts, const int* devices, const size_t* byte_starts) {
    if (int rc = ensure_init()) return rc;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    auto it = g_rt.allocs.find((char*)base);
    if (it == g_rt.allocs.end()) return fail(SCLX_E_INVALID_ARGUMENT, "array_repartition: unknown pointer");
    allocation* al = it->second;
    if (int rc = validate_parts(al->bytes, nparts, devices, byte_starts)) return rc;
    allocation probe;
    probe.va_bytes = al->va_bytes;
    build_runs(probe, nparts, devices, byte_starts);
    bool same = probe.runs.size() == al->runs.size();
    for (size_t i = 0; same && i < probe.runs.size(); ++i)
        same = probe.runs[i].offset == al->runs[i].offset && probe.runs[i].bytes == al->runs[i].bytes
            && physical_of(probe.runs[i].logical_device) == physical_of(al->runs[i].logical_device);
    if (same) {
        for (size_t i = 0; i < probe.runs.size(); ++i) al->runs[i].logical_device = probe.runs[i].logical_device;
        return 0;
    }
    if (int rc = flush_and_invalidate(*al)) return rc;
    if (int rc = sync_all_streams()) return rc;
    // stage the contents in a temporary range with the NEW placement, then swap mappings
    void* tmp = nullptr;
    if (int rc = sclx_b200_array_create(al->bytes, nparts, devices, byte_starts, &tmp)) return rc;
    allocation* nl = g_rt.allocs[(char*)tmp];
    if (int rc = sync_all_streams()) return rc;
    // piecewise: every piece has one source device and one destination device
    if (int rc = for_each_piece(*al, 0, al->bytes, [&](size_t o, size_t piece) -> int {
            return for_each_piece(*nl, o, piece, [&](size_t o2, size_t piece2) -> int {
                RT(cudaMemcpy(nl->base + o2, al->base + o2, piece2, cudaMemcpyDeviceToDevice));
                return 0;
            });
        }))
        return rc;
    if (int rc = sync_all_streams()) return rc;
    // unmap both, then map the new physical runs at the old address
    std::vector<allocation::run> new_runs = nl->runs;
    for (auto& r : nl->runs) CU(g_drv.MemUnmap((CUdeviceptr)(nl->base + r.offset), r.bytes));
    nl->runs.clear();
    if (int rc = unmap_runs(*al)) return rc;
    std::vector<CUmemAccessDesc> acc;
    const int nphys = g_rt.emulated ? 1 : g_rt.physical_count;
    for (int p = 0; p < nphys; ++p) {
        CUmemAccessDesc d{};
        d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        d.location.id   = p;
        d.flags         = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        acc.push_back(d);
    }
    for (auto& r : new_runs) {
        CU(g_drv.MemMap((CUdeviceptr)(al->base + r.offset), r.bytes, 0, r.handle, 0));
        al->runs.push_back(r);
    }
    CU(g_drv.MemSetAccess((CUdeviceptr)al->base, al->va_bytes, acc.data(), acc.size()));
    return sclx_b200_array_free(tmp);
}

// Arrays captured BY VALUE in a device lambda: scan the closure's object representation for
// 8-byte words that are the base address of a live runtime array.  The range dispatcher pulls
// the next grid-stride trip of such arrays into L2 while the current trip's loads are in
// flight (detail/b200/dispatch.cuh); a wrong guess costs a useless prefetch, never a fault.
int sclx_b200_rt_find_captured_arrays(const void* closure, size_t closure_bytes, int max_out, void** bases,
                                      size_t* bytes) {
    if (!closure || !bases || !bytes || max_out <= 0 || !g_rt.initialised) return 0;
    const char* p   = static_cast<const char*>(closure);
    const size_t lo = (8 - (reinterpret_cast<uintptr_t>(p) & 7)) & 7;
    int found       = 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    for (size_t off = lo; off + 8 <= closure_bytes && found < max_out; off += 8) {
        char* word = nullptr;
        memcpy(&word, p + off, 8);
        if (!word) continue;
        auto it = g_rt.allocs.find(word);
        if (it == g_rt.allocs.end()) continue;
        bool seen = false;
        for (int k = 0; k < found; ++k) seen |= bases[k] == word;
        if (seen) continue;
        bases[found] = word;
        bytes[found] = it->second->bytes;
        ++found;
    }
    return found;
}

int sclx_b200_array_owner(const void* ptr) {
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    allocation* al = find_alloc(ptr);
    if (!al) return -1;
    return owner_of(*al, size_t((char*)ptr - al->base));
}

// Called before ANY device-side access the runtime knows about (kernel launches, library
// kernels, device copies): host-written pages go back to the device and every host-valid page
// is dropped, so later host reads re-fetch what the device wrote.
int sclx_b200_rt_before_device_access(void) {
    if (!g_rt.initialised) return ensure_init();
    if (g_rt.host_valid_chunks.load() == 0) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    for (auto& kv : g_rt.allocs)
        if (int rc = flush_and_invalidate(*kv.second)) return rc;
    return 0;
}

// bulk "prefetch to the CPU" (array::prefetch_async({cpu_device_id})): the range becomes
// HOST-RESIDENT — fetched in large copies, readable AND writable without further faults
// (so read(2)/fread/recv into it work, as they do on the reference's managed memory,
// /root/reference scalix/array.cuh:241-248,647-711), and written back as a whole before the
// next device access, like migrated pages.  A prefetch of 64 MiB or more that covers at least a
// quarter of the array pins the mirror so both directions run at the PCIe rate.
int sclx_b200_rt_prefetch_to_host(const void* ptr, size_t bytes) {
    if (int rc = ensure_init()) return rc;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    allocation* al = find_alloc(ptr);
    if (!al || !al->overlay_ok || bytes == 0) return 0;
    size_t off = size_t((char*)ptr - al->base);
    size_t c0 = off / kMirrorChunk, c1 = (std::min(off + bytes, al->va_bytes) + kMirrorChunk - 1) / kMirrorChunk;
    if (!al->pinned && bytes >= kPinThreshold && bytes >= al->bytes / 4 && !getenv("SCLX_B200_NO_PIN")) {
        if (cudaHostRegister(al->alias, al->va_bytes, cudaHostRegisterPortable) == cudaSuccess) al->pinned = true;
        else cudaGetLastError();
    }
    return fetch_chunks(*al, c0, c1, /*for_write=*/true);
}

int sclx_b200_rt_quiesce(void) {
    if (int rc = ensure_init()) return rc;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    return sync_all_streams();
}

// ---- launch ordering ---------------------------------------------------------------------------
// A launch group = one kernel per participating logical device.  If earlier groups have not
// been synchronised by the host, every participating stream first waits for the previous
// group's events on all OTHER devices, so consecutive launches behave as if the caller had
// waited on each future (what the reference's users do), without a host round trip.
int sclx_b200_rt_begin_launch(int ndev, const int* devices) {
    if (int rc = ensure_init()) return rc;
    if (int rc = sclx_b200_rt_before_device_access()) return rc;
    if (!work_outstanding() || g_rt.dev.size() == 1) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    // events and stream waits do not need the stream's device to be current
    const unsigned long long synced = g_rt.synced_seq.load();
    std::vector<int> busy;  // devices with work the host has not waited for
    for (size_t o = 0; o < g_rt.dev.size(); ++o) {
        logical_device& ld = g_rt.dev[o];
        if (ld.last_launch_seq <= synced || ld.last_launch_seq <= ld.host_synced_seq) continue;
        if (ld.recorded_seq < ld.last_launch_seq) {
            RT(cudaEventRecord(ld.done, ld.stream));
            ld.recorded_seq = ld.last_launch_seq;
        }
        busy.push_back((int)o);
    }
    if (busy.empty()) return 0;
    auto waits_for_busy = [&](logical_device& me, int self) -> int {
        for (int o : busy)
            if (o != self) RT(cudaStreamWaitEvent(me.stream, g_rt.dev[o].done, 0));
        return 0;
    };
    if (ndev <= 2 || busy.size() <= 1) {
        for (int i = 0; i < ndev; ++i)
            if (int rc = waits_for_busy(g_rt.dev[devices[i]], devices[i])) return rc;
    } else {
        // O(N) instead of O(N^2) driver calls: the first participant (hub) waits for every
        // busy device, records its gate, the other participants wait for the gate
        logical_device& hub = g_rt.dev[devices[0]];
        if (int rc = waits_for_busy(hub, devices[0])) return rc;
        RT(cudaEventRecord(hub.gate, hub.stream));
        for (int i = 1; i < ndev; ++i) RT(cudaStreamWaitEvent(g_rt.dev[devices[i]].stream, hub.gate, 0));
    }
    return 0;
}

int sclx_b200_rt_end_launch(int ndev, const int* devices) {
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    const unsigned long long seq = g_rt.launch_seq.fetch_add(1) + 1;
    for (int i = 0; i < ndev; ++i)
        if (devices[i] >= 0 && devices[i] < (int)g_rt.dev.size()) g_rt.dev[devices[i]].last_launch_seq = seq;
    return 0;
}

// make stream `waiter` wait for everything issued so far on `signaller`'s stream
int sclx_b200_rt_stream_wait(int waiter, int signaller) {
    if (int rc = ensure_init()) return rc;
    if (waiter == signaller) return 0;
    if (waiter < 0 || signaller < 0 || waiter >= (int)g_rt.dev.size() || signaller >= (int)g_rt.dev.size())
        return fail(SCLX_E_INVALID_ARGUMENT, "stream_wait: device id out of range");
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    logical_device& s = g_rt.dev[signaller];
    logical_device& w = g_rt.dev[waiter];
    RT(cudaEventRecord(s.done, s.stream));
    s.recorded_seq = s.last_launch_seq;
    RT(cudaStreamWaitEvent(w.stream, s.done, 0));
    return 0;
}

// the two halves of sclx_b200_rt_stream_wait, for launch groups whose second kernel on device i
// depends on the FIRST kernel of other devices (chained groups, detail/execute_kernel.cuh): every
// device's issuing thread records its own event, then waits for the others'.  No lock: each
// call touches one device's event, and the callers order records before waits themselves.
int sclx_b200_rt_record_done(int logical) {
    if (logical < 0 || logical >= (int)g_rt.dev.size()) return fail(SCLX_E_INVALID_ARGUMENT, "record_done: bad device");
    RT(cudaEventRecord(g_rt.dev[logical].done, g_rt.dev[logical].stream));
    return 0;
}
int sclx_b200_rt_wait_done(int waiter, int signaller) {
    if (waiter < 0 || signaller < 0 || waiter >= (int)g_rt.dev.size() || signaller >= (int)g_rt.dev.size())
        return fail(SCLX_E_INVALID_ARGUMENT, "wait_done: bad device");
    if (waiter == signaller) return 0;
    RT(cudaStreamWaitEvent(g_rt.dev[waiter].stream, g_rt.dev[signaller].done, 0));
    return 0;
}

// queue the device's next completion ticket behind everything issued on its stream so far
int sclx_b200_rt_signal_done(int logical) {
    if (logical < 0 || logical >= (int)g_rt.dev.size()) return fail(SCLX_E_INVALID_ARGUMENT, "signal_done: bad device");
    if (logical >= kMaxSignalDevices || !g_signal[logical].word) return 0;  // waits use the stream itself
    device_signal& sg = g_signal[logical];
    std::lock_guard<std::mutex> lock(sg.mu);
    const unsigned ticket = sg.issued + 1;
    CUresult r = g_drv.StreamWriteValue32((CUstream)g_rt.dev[logical].stream, (CUdeviceptr)(uintptr_t)sg.word, ticket, 0);
    if (r != CUDA_SUCCESS) {
        // no stream memory operations here: fall back to stream synchronisation for good
        sg.word = nullptr;
        return 0;
    }
    sg.issued = ticket;
    return 0;
}

namespace {
// everything queued on `logical`'s stream by launch groups has finished (all of them end with a
// ticket); a stream error, or a stream that went idle without reaching the ticket, ends the
// poll with a real synchronisation, which reports it
cudaError_t wait_for_device(int logical) {
    logical_device& ld = g_rt.dev[logical];
    if (logical >= kMaxSignalDevices || !g_signal[logical].word) return cudaStreamSynchronize(ld.stream);
    device_signal& sg     = g_signal[logical];
    const unsigned target = sg.issued;
    unsigned long long spins = 0;
    while (int(*sg.word - target) < 0) {
        if ((++spins & 0x3fff) == 0) {
            cudaError_t e = cudaStreamQuery(ld.stream);
            if (e != cudaErrorNotReady) {
                if (e != cudaSuccess) return e;
                if (int(*sg.word - target) < 0) return cudaStreamSynchronize(ld.stream);
            }
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    return cudaSuccess;
}
}  // namespace

int sclx_b200_rt_sync_devices(int ndev, const int* devices) {
    if (int rc = ensure_init()) return rc;
    const int total = (int)g_rt.dev.size();
    if (ndev < 0 || (ndev > 0 && !devices)) return fail(SCLX_E_INVALID_ARGUMENT, "sync_devices: bad device list");
    unsigned long long seen = 0;  // bitmap of distinct logical devices (lists may repeat ids)
    for (int i = 0; i < ndev; ++i) {
        if (devices[i] < 0 || devices[i] >= total || devices[i] >= 64)
            return fail(SCLX_E_INVALID_ARGUMENT, "sync_devices: device id out of range");
        seen |= 1ull << devices[i];
    }
    const unsigned long long seq = g_rt.launch_seq.load();
    cudaError_t first = cudaSuccess;
    int distinct = 0;
    for (int d = 0; d < total && d < 64; ++d) {
        if (!(seen >> d & 1)) continue;
        ++distinct;
        cudaError_t e = wait_for_device(d);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    if (first != cudaSuccess) return rt_fail(first, "cudaStreamSynchronize");
    // only a sync of EVERY logical device covers all the groups issued before it started
    if (distinct == total) note_synced(seq);
    else {
        std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
        for (int d = 0; d < total && d < 64; ++d)
            if ((seen >> d & 1) && g_rt.dev[d].host_synced_seq < seq) g_rt.dev[d].host_synced_seq = seq;
    }
    return 0;
}

// ---- host-polled completion ------------------------------------------------------------------
// A kernel that KNOWS it is the last piece of a multi-device operation (the partition that
// arrives last in reduce_last_dim's fused fold) stores a ticket into a word of pinned, mapped
// host memory; the caller polls that word instead of synchronising N streams one after the
// other.  Flags are leased per call; the value is a process-wide ticket, so a stale flag can
// never satisfy a later wait.
namespace {
struct host_flag_pool {
    std::mutex mu;
    std::vector<unsigned*> free_flags;
    std::atomic<unsigned> ticket{0};
} g_flags;
}  // namespace

int sclx_b200_rt_host_flag_acquire(unsigned** flag, unsigned* ticket) {
    if (int rc = ensure_init()) return rc;
    if (!flag || !ticket) return fail(SCLX_E_INVALID_ARGUMENT, "host_flag_acquire: null out pointer");
    {
        std::lock_guard<std::mutex> lock(g_flags.mu);
        if (!g_flags.free_flags.empty()) {
            *flag = g_flags.free_flags.back();
            g_flags.free_flags.pop_back();
        } else {
            void* block = nullptr;  // 64 flags at a time, one cache line each
            RT(cudaHostAlloc(&block, 64 * 64, cudaHostAllocPortable | cudaHostAllocMapped));
            memset(block, 0, 64 * 64);
            for (int i = 1; i < 64; ++i) g_flags.free_flags.push_back(reinterpret_cast<unsigned*>((char*)block + 64 * i));
            *flag = reinterpret_cast<unsigned*>(block);
        }
    }
    unsigned t = g_flags.ticket.fetch_add(1) + 1;
    if (t == 0) t = g_flags.ticket.fetch_add(1) + 1;  // 0 is "never signalled"
    *ticket = t;
    return 0;
}

int sclx_b200_rt_host_flag_release(unsigned* flag) {
    if (!flag) return 0;
    std::lock_guard<std::mutex> lock(g_flags.mu);
    g_flags.free_flags.push_back(flag);
    return 0;
}

// Wait until *flag == ticket.  The devices' streams are checked now and then: a launch or
// execution error, or streams that went idle without the flag being set, end the wait with a
// real synchronisation (which reports the error).  On success the devices count as
// synchronised: everything the operation wrote was written before the flag.
int sclx_b200_rt_host_flag_wait(const unsigned* flag, unsigned ticket, int ndev, const int* devices) {
    if (int rc = ensure_init()) return rc;
    const volatile unsigned* f       = flag;
    const unsigned long long seq     = g_rt.launch_seq.load();
    const int total                  = (int)g_rt.dev.size();
    unsigned long long spins         = 0;
    while (*f != ticket) {
        if ((++spins & 0x3fff) == 0) {
            bool all_idle = true, failed = false;
            for (int i = 0; i < ndev; ++i) {
                if (devices[i] < 0 || devices[i] >= total) continue;
                cudaError_t e = cudaStreamQuery(g_rt.dev[devices[i]].stream);
                if (e == cudaErrorNotReady) all_idle = false;
                else if (e != cudaSuccess) failed = true;
            }
            if (failed || (all_idle && *f != ticket)) return sclx_b200_rt_sync_devices(ndev, devices);
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    unsigned long long seen = 0;
    for (int i = 0; i < ndev; ++i)
        if (devices[i] >= 0 && devices[i] < total && devices[i] < 64) seen |= 1ull << devices[i];
    int distinct = 0;
    for (int d = 0; d < total && d < 64; ++d) distinct += int(seen >> d & 1);
    if (distinct == total) note_synced(seq);
    else {
        std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
        for (int d = 0; d < total && d < 64; ++d)
            if ((seen >> d & 1) && g_rt.dev[d].host_synced_seq < seq) g_rt.dev[d].host_synced_seq = seq;
    }
    return 0;
}

// 1 when every stream in the list is idle, 0 when work is still running (never blocks)
int sclx_b200_rt_devices_idle(int ndev, const int* devices) {
    if (ensure_init()) return 0;
    const int total = (int)g_rt.dev.size();
    for (int i = 0; i < ndev; ++i) {
        const int d = devices[i];
        if (d < 0 || d >= total) continue;
        if (d < kMaxSignalDevices && g_signal[d].word) {
            if (int(*g_signal[d].word - g_signal[d].issued) < 0) return 0;
            continue;
        }
        // running, or failed: the blocking sync reports the error
        if (cudaStreamQuery(g_rt.dev[d].stream) != cudaSuccess) return 0;
    }
    return 1;
}

// ---- algorithm workspaces: leased per call ---------------------------------------------------
// A lease belongs to ONE algorithm call on ONE logical device; its kernels use it in stream
// order on that device's stream.  Two calls in flight (two host threads) never share partial
// buffers, tickets or look-back descriptors.
int sclx_b200_rt_workspace_acquire(int logical, void** lease) {
    if (int rc = ensure_init()) return rc;
    if (logical < 0 || logical >= (int)g_rt.dev.size() || !lease) return fail(SCLX_E_INVALID_ARGUMENT, "bad device");
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    logical_device& ld = g_rt.dev[logical];
    if (!ld.free_ws.empty()) {
        *lease = ld.free_ws.back();
        ld.free_ws.pop_back();
        return 0;
    }
    int cur = 0;
    cudaGetDevice(&cur);
    RT(cudaSetDevice(ld.physical));
    auto* ws    = new kernel_workspace;
    ws->logical = logical;
    cudaError_t e = cudaMalloc(&ws->ticket, 256);
    if (e == cudaSuccess) e = cudaMemset(ws->ticket, 0, 256);
    cudaSetDevice(cur);
    if (e != cudaSuccess) { delete ws; return rt_fail(e, "workspace allocation"); }
    *lease = ws;
    return 0;
}

// `poisoned`: the call ended in an exception and may have left tickets half-way: wait for the
// device and zero them again before the lease is reused
int sclx_b200_rt_workspace_release(void* lease, int poisoned) {
    if (!lease) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    auto* ws = static_cast<kernel_workspace*>(lease);
    if (poisoned && ws->ticket) {
        logical_device& ld = g_rt.dev[ws->logical];
        cudaStreamSynchronize(ld.stream);
        cudaMemcpy(ws->ticket, std::vector<unsigned char>(256, 0).data(), 256, cudaMemcpyHostToDevice);
        cudaGetLastError();
    }
    g_rt.dev[ws->logical].free_ws.push_back(ws);
    return 0;
}

// partials + ticket of reduce / reduce_last_dim / map_reduce
int sclx_b200_rt_workspace_reduce(void* lease, size_t bytes, void** partials, void** ticket) {
    if (!lease) return fail(SCLX_E_INVALID_ARGUMENT, "null workspace lease");
    auto* ws           = static_cast<kernel_workspace*>(lease);
    logical_device& ld = g_rt.dev[ws->logical];
    if (ws->partial_bytes < bytes) {
        int cur = 0;
        cudaGetDevice(&cur);
        RT(cudaSetDevice(ld.physical));
        // the holder's own earlier kernels may still read the old buffer
        if (ws->partials) { RT(cudaStreamSynchronize(ld.stream)); RT(cudaFree(ws->partials)); ws->partials = nullptr; }
        ws->partial_bytes = bytes + bytes / 2 + 4096;
        RT(cudaMalloc(&ws->partials, ws->partial_bytes));
        RT(cudaSetDevice(cur));
    }
    *partials = ws->partials;
    *ticket   = ws->ticket;
    return 0;
}

// look-back descriptors (`word_bytes` = 8 or 16, one buffer and epoch each) + tile-ticket
// counters of the scan kernels
int sclx_b200_rt_workspace_scan(void* lease, size_t tiles, int word_bytes, void** desc, unsigned* epoch,
                                void** counters) {
    if (!lease) return fail(SCLX_E_INVALID_ARGUMENT, "null workspace lease");
    if (word_bytes != 8 && word_bytes != 16) return fail(SCLX_E_INVALID_ARGUMENT, "descriptor words are 8 or 16 bytes");
    auto* ws           = static_cast<kernel_workspace*>(lease);
    logical_device& ld = g_rt.dev[ws->logical];
    const int k        = word_bytes == 8 ? 0 : 1;
    int cur = 0;
    cudaGetDevice(&cur);
    RT(cudaSetDevice(ld.physical));
    const size_t want = tiles * size_t(word_bytes) + 4096;
    if (ws->desc_bytes[k] < want) {
        if (ws->desc[k]) { RT(cudaStreamSynchronize(ld.stream)); RT(cudaFree(ws->desc[k])); ws->desc[k] = nullptr; }
        ws->desc_bytes[k] = want + want / 2;
        RT(cudaMalloc(&ws->desc[k], ws->desc_bytes[k]));
        RT(cudaMemsetAsync(ws->desc[k], 0, ws->desc_bytes[k], ld.stream));
        ws->epoch[k] = 0;
    }
    if (++ws->epoch[k] >= (1u << 30)) {
        RT(cudaMemsetAsync(ws->desc[k], 0, ws->desc_bytes[k], ld.stream));
        ws->epoch[k] = 1;
    }
    RT(cudaSetDevice(cur));
    *desc  = ws->desc[k];
    *epoch = ws->epoch[k];
    if (counters) *counters = ws->ticket + 32;
    return 0;
}

// small peer-visible scratch (per-partition totals, partial rows): VMM arrays on device 0,
// leased per call and returned when the call has synchronised
int sclx_b200_rt_scratch_acquire(size_t bytes, void** ptr) {
    if (int rc = ensure_init()) return rc;
    if (!ptr) return fail(SCLX_E_INVALID_ARGUMENT, "scratch: null out pointer");
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    for (auto& blk : g_rt.scratch)
        if (!blk.busy && blk.bytes >= bytes) {
            blk.busy = true;
            *ptr     = blk.ptr;
            return 0;
        }
    int dev0 = 0;
    size_t z = 0;
    const size_t want = std::max<size_t>(bytes, size_t(1) << 20);
    void* block = nullptr;
    if (int rc = sclx_b200_array_create(want, 1, &dev0, &z, &block)) return rc;
    g_rt.scratch.push_back({block, want, true});
    *ptr = block;
    return 0;
}

int sclx_b200_rt_scratch_release(void* ptr, int poisoned) {
    if (!ptr) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    for (auto& blk : g_rt.scratch)
        if (blk.ptr == ptr) {
            if (poisoned) {  // arrival counters may be half-way: zero the block once it is idle
                sync_all_streams();
                cudaMemset(blk.ptr, 0, blk.bytes);
                cudaDeviceSynchronize();
                cudaGetLastError();
            }
            blk.busy = false;
            return 0;
        }
    return fail(SCLX_E_INVALID_ARGUMENT, "scratch_release: unknown block");
}

// per-device, peer-visible, grow-only workspace of the exchange gather (request queues,
// response buffers, tile bookkeeping): a single-partition VMM array on `logical`
int sclx_b200_rt_exchange_workspace(int logical, size_t bytes, void** ptr) {
    if (int rc = ensure_init()) return rc;
    if (logical < 0 || logical >= (int)g_rt.dev.size()) return fail(SCLX_E_INVALID_ARGUMENT, "bad device");
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    logical_device& ld = g_rt.dev[logical];
    if (ld.exchange_bytes < bytes) {
        if (ld.exchange) {
            sync_all_streams();
            sclx_b200_array_free(ld.exchange);
            ld.exchange = nullptr;
            ld.exchange_bytes = 0;
        }
        size_t z = 0;
        size_t want = bytes + bytes / 8 + (size_t(1) << 21);
        if (int rc = sclx_b200_array_create(want, 1, &logical, &z, &ld.exchange)) return rc;
        ld.exchange_bytes = want;
    }
    *ptr = ld.exchange;
    return 0;
}

unsigned long long sclx_b200_rt_host_faults(void) { return g_rt.faults; }

// array pool: bytes parked / reuse counters / explicit release (e.g. before a large cudaMalloc)
size_t sclx_b200_rt_pool_bytes(void) { return g_rt.pool_bytes; }
// device memory of `logical`'s GPU held by parked arrays (query_memory_status adds it to `free`)
size_t sclx_b200_rt_pool_bytes_on(int logical) {
    if (!g_rt.initialised || logical < 0 || logical >= (int)g_rt.dev.size()) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    size_t sum = 0;
    for (allocation* al : g_rt.pool)
        for (auto& r : al->runs)
            if (physical_of(r.logical_device) == physical_of(logical)) sum += r.bytes;
    return sum;
}
unsigned long long sclx_b200_rt_pool_hits(void) { return g_rt.pool_hits; }
int sclx_b200_rt_pool_release(void) {
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    pool_purge(0);
    return 0;
}

}  // extern "C"
