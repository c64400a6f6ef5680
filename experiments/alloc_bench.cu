// Experiment: cost of sclx::array construction / destruction (VMM create+map+access+zero+mirror)
#include <scalix/array.cuh>
#include <scalix/fill.cuh>
#include <chrono>
#include <cstdio>
int main() {
    using clk = std::chrono::steady_clock;
    { sclx::array<float, 1> warm({1024}); }
    for (size_t mb : {1, 16, 256, 1024, 4096}) {
        const size_t n = mb * (size_t(1) << 20) / 4;
        double create = 0, destroy = 0;
        const int reps = 3;
        for (int r = 0; r < reps; ++r) {
            auto t0 = clk::now();
            auto* a = new sclx::array<float, 1>({n});
            cudaDeviceSynchronize();
            auto t1 = clk::now();
            delete a;
            auto t2 = clk::now();
            create += std::chrono::duration<double, std::milli>(t1 - t0).count();
            destroy += std::chrono::duration<double, std::milli>(t2 - t1).count();
        }
        std::printf("%6zu MiB  create %9.3f ms  destroy %9.3f ms\n", mb, create / reps, destroy / reps);
    }
    return 0;
}
