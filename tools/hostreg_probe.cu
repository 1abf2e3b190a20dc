// cost of page-locking a pageable array in place (cudaHostRegister / cudaHostUnregister) against
// repacking it into page-locked staging with host threads.  Build: nvcc -O2 -o tools/_bin/hostreg_probe tools/hostreg_probe.cu
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <cuda_runtime.h>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main()
{
    cudaFree(0);
    for (size_t mb : {16, 122, 800}) {
        const size_t bytes = mb << 20;
        char *p = (char *)malloc(bytes);
        memset(p, 1, bytes);
        double best_r = 1e9, best_u = 1e9;
        for (int rep = 0; rep < 4; ++rep) {
            double t0 = now();
            cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterDefault);
            double t1 = now();
            if (e != cudaSuccess) { printf("register failed: %s\n", cudaGetErrorString(e)); return 1; }
            cudaHostUnregister(p);
            double t2 = now();
            best_r = std::min(best_r, t1 - t0);
            best_u = std::min(best_u, t2 - t1);
        }
        // threads copying into page-locked memory
        char *q;
        cudaMallocHost(&q, bytes);
        for (int nt : {4, 8, 16}) {
            double best = 1e9;
            for (int rep = 0; rep < 4; ++rep) {
                double t0 = now();
                std::vector<std::thread> th;
                for (int t = 0; t < nt; ++t)
                    th.emplace_back([=] { memcpy(q + bytes * t / nt, p + bytes * t / nt, bytes * (t + 1) / nt - bytes * t / nt); });
                for (auto &x : th) x.join();
                best = std::min(best, now() - t0);
            }
            printf("{\"MiB\": %zu, \"threads\": %d, \"copy_to_pinned_ms\": %.3f, \"GBps\": %.1f}\n", mb, nt, best * 1e3, bytes / best / 1e9);
        }
        printf("{\"MiB\": %zu, \"register_ms\": %.3f, \"unregister_ms\": %.3f}\n", mb, best_r * 1e3, best_u * 1e3);
        cudaFreeHost(q);
        free(p);
    }
    return 0;
}
