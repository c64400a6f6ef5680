// standalone probe of the C ABI (no python): prints each step unbuffered
#include "scalix_b200.h"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <execinfo.h>
#include <unistd.h>
static bool g_done=false;
static void at_exit_bt(){ if(g_done) return; void* bt[64]; int n=backtrace(bt,64); fprintf(stderr,"!! exit() called before PROBE_DONE, backtrace:\n"); backtrace_symbols_fd(bt,n,2);}
#define STEP(x) do { fprintf(stderr, "-> %s\n", #x); int rc = (x); fprintf(stderr, "   rc=%d %s\n", rc, rc ? sclx_b200_last_error() : ""); } while (0)
int main() {
    setvbuf(stderr, nullptr, _IONBF, 0);
    atexit(at_exit_bt);
    fprintf(stderr, "devices %d\n", sclx_b200_device_count());
    STEP(sclx_b200_set_device(0));
    void *a = nullptr, *b = nullptr, *t = nullptr;
    size_t n = 1 << 20;
    STEP(sclx_b200_malloc(&a, n * 8));
    STEP(sclx_b200_malloc(&b, n * 8));
    STEP(sclx_b200_malloc(&t, 64));
    int start = 1;
    STEP(sclx_b200_iota(a, n, SCLX_I32, &start, 0, nullptr));
    STEP(sclx_b200_stream_synchronize(nullptr));
    STEP(sclx_b200_inclusive_scan(a, b, 1, SCLX_I32, SCLX_PLUS, nullptr, 0, nullptr, nullptr));
    STEP(sclx_b200_stream_synchronize(nullptr));
    STEP(sclx_b200_inclusive_scan(a, b, n, SCLX_I32, SCLX_PLUS, nullptr, 0, t, nullptr));
    STEP(sclx_b200_stream_synchronize(nullptr));
    std::vector<int> h(n);
    STEP(sclx_b200_memcpy_d2h(h.data(), b, n * 4, nullptr));
    STEP(sclx_b200_stream_synchronize(nullptr));
    fprintf(stderr, "scan[0]=%d scan[n-1]=%d (want %lld)\n", h[0], h[n - 1], (long long)((n * (n + 1) / 2) & 0xffffffff));
    float ident = 0.f;
    STEP(sclx_b200_reduce_last_dim(a, 64, n / 64, SCLX_F32, SCLX_PLUS, &ident, b, nullptr));
    STEP(sclx_b200_stream_synchronize(nullptr));
    long long id64 = 0;
    STEP(sclx_b200_reduce(a, n / 2, SCLX_I64, SCLX_PLUS, &id64, t, nullptr));
    STEP(sclx_b200_stream_synchronize(nullptr));
    fprintf(stderr, "PROBE_DONE launches=%llu\n", sclx_b200_launch_count());
    g_done = true;
    return 0;
}
