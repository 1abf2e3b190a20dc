// How fast can host threads repack a pageable (rows x cols) array into a page-locked-like buffer?
// (Design evidence for the staging in csrc/mpx_api.cu; plain malloc on both sides.)
// Build: g++ -O2 -pthread -o tools/_bin/host_copy_probe tools/host_copy_probe.cpp
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static void copy_rows(double *dst, size_t dst_ld, const double *src, size_t src_ld, size_t cols, size_t rows, int nt)
{
    auto part = [=](size_t r0, size_t r1) {
        if (dst_ld == cols && src_ld == cols) memcpy(dst + r0 * cols, src + r0 * cols, (r1 - r0) * cols * 8);
        else for (size_t r = r0; r < r1; ++r) memcpy(dst + r * dst_ld, src + r * src_ld, cols * 8);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(part, rows * t / nt, rows * (t + 1) / nt);
    part(0, rows / nt);
    for (auto &x : th) x.join();
}

int main()
{
    const size_t N = 59711, k = 256;
    double *src = (double *)malloc(N * k * 8), *dst = (double *)malloc(N * k * 8);
    for (size_t i = 0; i < N * k; ++i) { src[i] = (double)i; dst[i] = 0; }
    for (int nt : {1, 2, 4, 8, 12, 16}) {
        for (int mode = 0; mode < 3; ++mode) {
            // 0: whole array contiguous; 1: 8 MiB row blocks; 2: four blocks of 64 columns
            double best = 1e9;
            for (int rep = 0; rep < 5; ++rep) {
                auto t0 = std::chrono::steady_clock::now();
                if (mode == 0) copy_rows(dst, k, src, k, k, N, nt);
                else if (mode == 1) {
                    const size_t rb = (8u << 20) / (k * 8);
                    for (size_t r0 = 0; r0 < N; r0 += rb) copy_rows(dst + r0 * k, k, src + r0 * k, k, k, std::min(rb, N - r0), nt);
                } else {
                    for (size_t c0 = 0; c0 < k; c0 += 64) copy_rows(dst + c0 * N, 64, src + c0, k, 64, N, nt);
                }
                double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                if (dt < best) best = dt;
            }
            printf("{\"threads\": %d, \"mode\": \"%s\", \"ms\": %.3f, \"GBps\": %.1f}\n", nt,
                   mode == 0 ? "contiguous" : mode == 1 ? "row blocks 8 MiB" : "column blocks 64", best * 1e3, N * k * 8 / best / 1e9);
        }
    }
    return 0;
}
