// Experiment (not product): can a host mapping be overlaid on a CUDA VMM virtual range so the
// same numeric pointer is dereferenceable on the host (software-managed mirror) while the GPU
// keeps seeing device memory?  Needed to keep the reference's raw-pointer host idioms
// (`result.data().get()[i]`, range-for over begin()/end()) working without Unified Memory.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <csignal>
#include <sys/mman.h>
#include <unistd.h>

#define CK(x) do { CUresult r = (x); if (r != CUDA_SUCCESS) { const char* s; cuGetErrorString(r, &s); printf("FAIL %s -> %s\n", #x, s); return 1; } } while (0)
#define CKR(x) do { cudaError_t r = (x); if (r != cudaSuccess) { printf("FAIL %s -> %s\n", #x, cudaGetErrorString(r)); return 1; } } while (0)

__global__ void write_pattern(unsigned* p, size_t n, unsigned base) {
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) p[i] = base + unsigned(i);
}
__global__ void check_pattern(const unsigned* p, size_t n, unsigned base, unsigned* bad) {
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        if (p[i] != base + unsigned(i)) atomicAdd(bad, 1u);
}

static char* g_base; static size_t g_size; static void* g_bounce; static int g_faults;
static void on_segv(int, siginfo_t* si, void*) {
    char* a = (char*)si->si_addr;
    if (a < g_base || a >= g_base + g_size) { _exit(99); }
    size_t page = 2u << 20;
    char* pg = g_base + ((a - g_base) / page) * page;
    mprotect(pg, page, PROT_READ | PROT_WRITE);
    cudaMemcpy(g_bounce, pg, page, cudaMemcpyDeviceToHost);   // pg resolves to DEVICE memory in UVA
    memcpy(pg, g_bounce, page);                                 // CPU store goes to the host overlay
    ++g_faults;
}

int main() {
    CKR(cudaSetDevice(0)); CKR(cudaFree(0));
    size_t gran = 0;
    CUmemAllocationProp prop{}; prop.type = CU_MEM_ALLOCATION_TYPE_PINNED; prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE; prop.location.id = 0;
    CK(cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM));
    printf("granularity %zu\n", gran);
    size_t size = 64u << 20; size_t n = size / 4;
    CUdeviceptr va = 0;
    CK(cuMemAddressReserve(&va, size, 0, 0, 0));
    CUmemGenericAllocationHandle h;
    CK(cuMemCreate(&h, size, &prop, 0));
    CK(cuMemMap(va, size, 0, h, 0));
    CUmemAccessDesc acc{}; acc.location = prop.location; acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CK(cuMemSetAccess(va, size, &acc, 1));
    printf("va %p\n", (void*)va);
    write_pattern<<<148, 256>>>((unsigned*)va, n, 7u);
    CKR(cudaDeviceSynchronize());
    // how did the driver reserve the range on the host side?
    { char cmd[256]; snprintf(cmd, sizeof cmd, "grep -i '^%lx' /proc/%d/maps | head -3", (unsigned long)va, getpid()); fflush(stdout); if (system(cmd)) printf("(range not at start of a maps entry)\n"); }
    // overlay host pages at the same VA
    void* m = mmap((void*)va, size, PROT_NONE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_FIXED | MAP_NORESERVE, -1, 0);
    printf("overlay mmap -> %p (%s)\n", m, m == (void*)va ? "ok" : "FAILED");
    if (m != (void*)va) return 2;
    // GPU must still see device memory
    unsigned* bad; CKR(cudaMalloc(&bad, 4)); CKR(cudaMemset(bad, 0, 4));
    check_pattern<<<148, 256>>>((const unsigned*)va, n, 7u, bad);
    CKR(cudaDeviceSynchronize());
    unsigned hb = 1; CKR(cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost));
    printf("GPU view after overlay: %u mismatches\n", hb);
    // cudaMemcpy must still treat the pointer as device memory
    CKR(cudaMallocHost(&g_bounce, 2u << 20));
    CKR(cudaMemcpy(g_bounce, (void*)va, 4096, cudaMemcpyDeviceToHost));
    printf("cudaMemcpy D2H first word %u (want 7)\n", ((unsigned*)g_bounce)[0]);
    // demand-fill through SIGSEGV
    g_base = (char*)va; g_size = size;
    struct sigaction sa{}; sa.sa_sigaction = on_segv; sa.sa_flags = SA_SIGINFO; sigaction(SIGSEGV, &sa, nullptr);
    unsigned* hp = (unsigned*)va;
    unsigned long long sum = 0; size_t wrong = 0;
    for (size_t i = 0; i < n; i += 1021) { if (hp[i] != 7u + unsigned(i)) ++wrong; sum += hp[i]; }
    printf("host reads through overlay: faults=%d wrong=%zu\n", g_faults, wrong);
    // host write -> flush back to the device -> GPU sees it
    hp[12345] = 0xC0FFEEu;
    memcpy(g_bounce, (char*)va, 2u << 20);
    CKR(cudaMemcpy((void*)va, g_bounce, 2u << 20, cudaMemcpyHostToDevice));
    unsigned v = 0; CKR(cudaMemcpy(&v, (char*)va + 12345 * 4, 4, cudaMemcpyDeviceToHost));
    printf("device sees host write: %x (want c0ffee)\n", v);
    // thrust-style: raw pointer arithmetic + kernel on the same numeric pointer still fine
    write_pattern<<<148, 256>>>((unsigned*)va, n, 100u);
    CKR(cudaDeviceSynchronize());
    CKR(cudaMemcpy(&v, (char*)va + 4, 4, cudaMemcpyDeviceToHost));
    printf("after second kernel word1=%u (want 101)\n", v);
    CK(cuMemUnmap(va, size)); CK(cuMemRelease(h)); CK(cuMemAddressFree(va, size));
    printf("OVERLAY_EXPERIMENT_DONE\n");
    return 0;
}
