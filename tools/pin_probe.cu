// Dev tool: how fast can this box produce page-locked memory?  (cold-start budget of the streaming writer)
//   nvcc -O2 -o /tmp/pin_probe tools/pin_probe.cu && /tmp/pin_probe
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
    const size_t G = 1ull << 30;
    double t0 = now();
    cudaFree(0);
    printf("cuda init + primary context: %.3f s\n", now() - t0);
    void *p;
    t0 = now(); cudaHostAlloc(&p, G, cudaHostAllocPortable); double t = now() - t0;
    printf("cudaHostAlloc 1 GiB: %.3f s (%.2f GB/s)\n", t, G / t / 1e9);
    t0 = now(); cudaFreeHost(p); printf("cudaFreeHost: %.3f s\n", now() - t0);
    t0 = now(); cudaHostAlloc(&p, G, cudaHostAllocPortable); t = now() - t0;
    printf("cudaHostAlloc 1 GiB again: %.3f s (%.2f GB/s)\n", t, G / t / 1e9);
    cudaFreeHost(p);
    {   // 8 threads x 128 MiB
        std::vector<std::thread> th; void *q[8];
        t0 = now();
        for (int i = 0; i < 8; ++i) th.emplace_back([&, i] { cudaHostAlloc(&q[i], G / 8, cudaHostAllocPortable); });
        for (auto &x : th) x.join();
        t = now() - t0;
        printf("8 threads x cudaHostAlloc 128 MiB: %.3f s (%.2f GB/s)\n", t, G / t / 1e9);
        for (int i = 0; i < 8; ++i) cudaFreeHost(q[i]);
    }
    for (int huge = 0; huge < 2; ++huge) {
        void *m = mmap(nullptr, G, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (huge) madvise(m, G, MADV_HUGEPAGE);
        t0 = now();
        {   // touch with 8 threads
            std::vector<std::thread> th;
            for (int i = 0; i < 8; ++i) th.emplace_back([&, i] { memset((char *)m + i * (G / 8), 1, G / 8); });
            for (auto &x : th) x.join();
        }
        double tt = now() - t0;
        t0 = now(); cudaError_t e = cudaHostRegister(m, G, cudaHostRegisterPortable); t = now() - t0;
        printf("mmap%s: touch (8 thr) %.3f s, cudaHostRegister %.3f s (%s) -> %.2f GB/s total\n", huge ? "+MADV_HUGEPAGE" : "", tt, t,
               cudaGetErrorString(e), G / (t + tt) / 1e9);
        if (e == cudaSuccess) { t0 = now(); cudaHostUnregister(m); printf("   unregister %.3f s\n", now() - t0); }
        munmap(m, G);
    }
    {   // register in 64 MiB pieces on 4 threads
        void *m = mmap(nullptr, G, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        madvise(m, G, MADV_HUGEPAGE);
        memset(m, 1, G);
        t0 = now();
        std::vector<std::thread> th;
        for (int i = 0; i < 4; ++i) th.emplace_back([&, i] { for (int k = 0; k < 4; ++k) cudaHostRegister((char *)m + (i * 4 + k) * (G / 16), G / 16, cudaHostRegisterPortable); });
        for (auto &x : th) x.join();
        t = now() - t0;
        printf("register 16 x 64 MiB on 4 threads: %.3f s (%.2f GB/s)\n", t, G / t / 1e9);
    }
    return 0;
}
