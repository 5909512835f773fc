// Host-side cost of one kernel launch on this box, as the decoder issues them: cudaLaunchKernelEx
// with / without the programmatic-stream-serialization attribute, with small / decoder-sized
// by-value parameters, from 1 or 8 threads (one stream each) of one process.
// build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a tools/micro/launch_cost.cu -o /tmp/launch_cost -lpthread
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>

struct Small { int a[4]; };
struct Big { int a[200]; }; // 800 B: ScanGeom + ScanBufs + destination + quantisers

__global__ void k_small(Small s) { if (s.a[0] == 12345) printf("x"); }
__global__ void k_big(Big s) { if (s.a[0] == 12345) printf("x"); }
__global__ void k_big_pdl(Big s)
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (s.a[0] == 12345) printf("x");
}

template <class K, class P>
double run(K kernel, P p, bool pdl, cudaStream_t st, int n)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148);
    cfg.blockDim = dim3(128);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    if (pdl) { cfg.attrs = attr; cfg.numAttrs = 1; }
    for (int i = 0; i < 50; ++i) cudaLaunchKernelEx(&cfg, kernel, p);
    cudaStreamSynchronize(st);
    double tot = 0;
    for (int rep = 0; rep < n / 50; ++rep) { // bursts of 50 so that the queue never fills
        const auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < 50; ++i) cudaLaunchKernelEx(&cfg, kernel, p);
        tot += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
        cudaStreamSynchronize(st);
    }
    return tot / (n / 50 * 50);
}

int main()
{
    cudaFree(0);
    Small s{}; Big b{};
    for (int threads : {1, 8}) {
        std::vector<double> r(4 * threads);
        std::vector<std::thread> th;
        for (int t = 0; t < threads; ++t)
            th.emplace_back([&, t] {
                cudaStream_t st;
                cudaStreamCreate(&st);
                r[4 * t + 0] = run(k_small, s, false, st, 2000);
                r[4 * t + 1] = run(k_big, b, false, st, 2000);
                r[4 * t + 2] = run(k_big_pdl, b, true, st, 2000);
                r[4 * t + 3] = run(k_small, s, true, st, 2000);
                cudaStreamDestroy(st);
            });
        for (auto& x : th) x.join();
        double a[4] = {0, 0, 0, 0};
        for (int t = 0; t < threads; ++t)
            for (int i = 0; i < 4; ++i) a[i] += r[4 * t + i] / threads;
        printf("threads=%d  us/launch: small %.2f | 800B %.2f | 800B+PDL %.2f | small+PDL-attr %.2f\n", threads, a[0], a[1], a[2], a[3]);
    }
    return 0;
}
