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
};

struct logical_device {
    int physical        = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t done    = nullptr;  // recorded after the latest launch group
    cudaEvent_t gate    = nullptr;  // recorded by the hub of a launch group once every device's `done` has fired
    int sm_count        = 0;
    // algorithm workspaces
    void* scan_desc = nullptr; size_t scan_bytes = 0; unsigned scan_epoch = 0;
    void* partials = nullptr; size_t partial_bytes = 0; unsigned* ticket = nullptr;
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
    std::atomic<bool> unsynced{false};  // launches issued since the last full sync
    bool handler_installed = false;
    struct sigaction old_action {};
    void* scratch = nullptr; size_t scratch_bytes = 0;
    unsigned long long faults = 0, flushed_chunks = 0;
    // freed arrays kept mapped for reuse (VA + physical pages + peer access + mirror intact):
    // creating a multi-GiB peer-mapped range costs tens of ms on a multi-GPU node, reusing one
    // costs its zero-fill.  Oldest first; bounded by pool_cap bytes (SCLX_B200_POOL_BYTES).
    std::vector<allocation*> pool;
    size_t pool_bytes = 0, pool_cap = size_t(32) << 30;
    unsigned long long pool_hits = 0, pool_misses = 0;
};
runtime g_rt;

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
        RT(cudaMalloc(&ld.ticket, 256));
        RT(cudaMemset(ld.ticket, 0, 256));
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
int sync_all_streams() {
    int cur = 0;
    cudaGetDevice(&cur);
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
    g_rt.unsynced.store(false);
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

// bring chunks [c0, c1) of the host mirror up to date (device -> host)
int fetch_chunks(allocation& al, size_t c0, size_t c1, bool for_write) {
    // coalesce invalid chunks into runs so a bulk prefetch is one copy per physical run
    size_t c = c0;
    bool synced = false;
    while (c < c1) {
        if (al.state[c] != kInvalid) {
            if (for_write && al.state[c] == kClean) {
                al.state[c] = kDirty;
                mprotect(al.base + c * kMirrorChunk, kMirrorChunk, PROT_READ | PROT_WRITE);
            }
            ++c;
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
        if (int rc = for_each_piece(al, off, len, [&](size_t o, size_t piece) -> int {
                RT(cudaMemcpy(al.alias + o, al.base + o, piece, cudaMemcpyDeviceToHost));
                return 0;
            }))
            return rc;
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
            if (int rc = for_each_piece(al, off, len, [&](size_t o, size_t piece) -> int {
                    RT(cudaMemcpy(al.base + o, al.alias + o, piece, cudaMemcpyHostToDevice));
                    return 0;
                }))
                return rc;
            g_rt.flushed_chunks += f - d;
            d = f;
        }
        size_t off = c * kMirrorChunk, len = (e - c) * kMirrorChunk;
        mprotect(al.base + off, len, PROT_NONE);
        if (al.memfd >= 0) fallocate(al.memfd, FALLOC_FL_PUNCH_HOLE | FALLOC_FL_KEEP_SIZE, off, len);
        for (size_t k = c; k < e; ++k) al.state[k] = kInvalid;
        g_rt.host_valid_chunks.fetch_sub(e - c);
        c = e;
    }
    al.n_valid = 0;
    return 0;
}

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
            size_t c = size_t(addr - al->base) / kMirrorChunk;
            ++g_rt.faults;
            if (fetch_chunks(*al, c, c + 1, write) == 0) handled = true;
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
int destroy_allocation(allocation* al) {
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
    if (g_rt.unsynced.load()) sync_all_streams();  // kernels in flight may still use it
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

// bulk "prefetch to the CPU" (array::prefetch_async({cpu_device_id})): one large copy
int sclx_b200_rt_prefetch_to_host(const void* ptr, size_t bytes) {
    if (int rc = ensure_init()) return rc;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    allocation* al = find_alloc(ptr);
    if (!al || !al->overlay_ok || bytes == 0) return 0;
    size_t off = size_t((char*)ptr - al->base);
    size_t c0 = off / kMirrorChunk, c1 = (std::min(off + bytes, al->va_bytes) + kMirrorChunk - 1) / kMirrorChunk;
    return fetch_chunks(*al, c0, c1, false);
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
    if (!g_rt.unsynced.load() || g_rt.dev.size() == 1) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    int cur = 0;
    cudaGetDevice(&cur);
    if (ndev <= 2) {
        for (int i = 0; i < ndev; ++i) {
            logical_device& me = g_rt.dev[devices[i]];
            RT(cudaSetDevice(me.physical));
            for (size_t o = 0; o < g_rt.dev.size(); ++o)
                if ((int)o != devices[i]) RT(cudaStreamWaitEvent(me.stream, g_rt.dev[o].done, 0));
        }
    } else {
        // O(N) instead of O(N^2) driver calls: the first participant (hub) waits for every
        // other device, records its gate, the other participants wait for the gate
        logical_device& hub = g_rt.dev[devices[0]];
        RT(cudaSetDevice(hub.physical));
        for (size_t o = 0; o < g_rt.dev.size(); ++o)
            if ((int)o != devices[0]) RT(cudaStreamWaitEvent(hub.stream, g_rt.dev[o].done, 0));
        RT(cudaEventRecord(hub.gate, hub.stream));
        for (int i = 1; i < ndev; ++i) {
            logical_device& me = g_rt.dev[devices[i]];
            RT(cudaSetDevice(me.physical));
            RT(cudaStreamWaitEvent(me.stream, hub.gate, 0));
        }
    }
    RT(cudaSetDevice(cur));
    return 0;
}

int sclx_b200_rt_end_launch(int ndev, const int* devices) {
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    if (g_rt.dev.size() > 1) {
        int cur = 0;
        cudaGetDevice(&cur);
        for (int i = 0; i < ndev; ++i) {
            logical_device& me = g_rt.dev[devices[i]];
            RT(cudaSetDevice(me.physical));
            RT(cudaEventRecord(me.done, me.stream));
        }
        RT(cudaSetDevice(cur));
    }
    g_rt.unsynced.store(true);
    return 0;
}

// make stream `waiter` wait for everything issued so far on `signaller`'s stream
int sclx_b200_rt_stream_wait(int waiter, int signaller) {
    if (int rc = ensure_init()) return rc;
    if (waiter == signaller) return 0;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    int cur = 0;
    cudaGetDevice(&cur);
    logical_device& s = g_rt.dev[signaller];
    logical_device& w = g_rt.dev[waiter];
    RT(cudaSetDevice(s.physical));
    RT(cudaEventRecord(s.done, s.stream));
    RT(cudaSetDevice(w.physical));
    RT(cudaStreamWaitEvent(w.stream, s.done, 0));
    RT(cudaSetDevice(cur));
    return 0;
}

int sclx_b200_rt_sync_devices(int ndev, const int* devices) {
    if (int rc = ensure_init()) return rc;
    int cur = 0;
    cudaGetDevice(&cur);
    cudaError_t first = cudaSuccess;
    for (int i = 0; i < ndev; ++i) {
        logical_device& me = g_rt.dev[devices[i]];
        cudaSetDevice(me.physical);
        cudaError_t e = cudaStreamSynchronize(me.stream);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    cudaSetDevice(cur);
    if (first != cudaSuccess) return rt_fail(first, "cudaStreamSynchronize");
    if (ndev == (int)g_rt.dev.size()) g_rt.unsynced.store(false);
    return 0;
}

// ---- algorithm workspaces (per logical device; used in stream order on that device's stream) --
int sclx_b200_rt_scan_workspace(int logical, size_t tiles, void** desc, unsigned* epoch) {
    if (int rc = ensure_init()) return rc;
    if (logical < 0 || logical >= (int)g_rt.dev.size()) return fail(SCLX_E_INVALID_ARGUMENT, "bad device");
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    logical_device& ld = g_rt.dev[logical];
    int cur = 0;
    cudaGetDevice(&cur);
    RT(cudaSetDevice(ld.physical));
    const size_t want = tiles * 16 + 4096;
    if (ld.scan_bytes < want) {
        if (ld.scan_desc) { RT(cudaStreamSynchronize(ld.stream)); RT(cudaFree(ld.scan_desc)); }
        ld.scan_bytes = want + want / 2;
        RT(cudaMalloc(&ld.scan_desc, ld.scan_bytes));
        RT(cudaMemsetAsync(ld.scan_desc, 0, ld.scan_bytes, ld.stream));
        ld.scan_epoch = 0;
    }
    if (++ld.scan_epoch >= (1u << 30)) {
        RT(cudaMemsetAsync(ld.scan_desc, 0, ld.scan_bytes, ld.stream));
        ld.scan_epoch = 1;
    }
    RT(cudaSetDevice(cur));
    *desc  = ld.scan_desc;
    *epoch = ld.scan_epoch;
    return 0;
}

// tile-ticket counters of the streaming scan (zero-initialised, self-resetting)
int sclx_b200_rt_scan_counters(int logical, void** counters) {
    if (int rc = ensure_init()) return rc;
    if (logical < 0 || logical >= (int)g_rt.dev.size() || !counters)
        return fail(SCLX_E_INVALID_ARGUMENT, "bad device");
    *counters = g_rt.dev[logical].ticket + 32;
    return 0;
}

int sclx_b200_rt_reduce_workspace(int logical, size_t bytes, void** partials, void** ticket) {
    if (int rc = ensure_init()) return rc;
    if (logical < 0 || logical >= (int)g_rt.dev.size()) return fail(SCLX_E_INVALID_ARGUMENT, "bad device");
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    logical_device& ld = g_rt.dev[logical];
    int cur = 0;
    cudaGetDevice(&cur);
    RT(cudaSetDevice(ld.physical));
    if (ld.partial_bytes < bytes) {
        if (ld.partials) { RT(cudaStreamSynchronize(ld.stream)); RT(cudaFree(ld.partials)); }
        ld.partial_bytes = bytes + bytes / 2 + 4096;
        RT(cudaMalloc(&ld.partials, ld.partial_bytes));
    }
    RT(cudaSetDevice(cur));
    *partials = ld.partials;
    *ticket   = ld.ticket;
    return 0;
}

// small peer-visible scratch (per-partition totals, partial rows): a VMM array on device 0
int sclx_b200_rt_scratch(size_t bytes, void** ptr) {
    if (int rc = ensure_init()) return rc;
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    if (g_rt.scratch_bytes < bytes) {
        if (g_rt.scratch) {
            sync_all_streams();
            sclx_b200_array_free(g_rt.scratch);
            g_rt.scratch = nullptr;
        }
        int dev0 = 0;
        size_t z = 0;
        size_t want = std::max<size_t>(bytes, size_t(1) << 20);
        if (int rc = sclx_b200_array_create(want, 1, &dev0, &z, &g_rt.scratch)) return rc;
        g_rt.scratch_bytes = want;
    }
    *ptr = g_rt.scratch;
    return 0;
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
unsigned long long sclx_b200_rt_pool_hits(void) { return g_rt.pool_hits; }
int sclx_b200_rt_pool_release(void) {
    std::lock_guard<std::recursive_mutex> lock(g_rt.mu);
    pool_purge(0);
    return 0;
}

}  // extern "C"
