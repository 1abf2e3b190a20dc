// PCIe copy rates for the layouts mpx_dds_host can choose from: contiguous vs 2-D (row pieces of a
// pitched array) in both directions, page-locked host memory.  Build: nvcc -O2 -o tools/_bin/copy2d_probe tools/copy2d_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

static float timeit(cudaStream_t st, int reps, void (*fn)(void *), void *ctx)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    fn(ctx);
    cudaStreamSynchronize(st);
    float best = 1e9f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(a, st);
        fn(ctx);
        cudaEventRecord(b, st);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

struct Ctx { double *h, *d; size_t N, k, kc; cudaStream_t st; int mode; };

static void run(void *p)
{
    Ctx &c = *(Ctx *)p;
    const size_t nb = c.k / c.kc;
    switch (c.mode) {
    case 0: cudaMemcpyAsync(c.d, c.h, c.N * c.k * 8, cudaMemcpyHostToDevice, c.st); break;
    case 1: cudaMemcpyAsync(c.h, c.d, c.N * c.k * 8, cudaMemcpyDeviceToHost, c.st); break;
    case 2:   // H2D, host dense blocks -> device pitched columns
        for (size_t j = 0; j < nb; ++j)
            cudaMemcpy2DAsync(c.d + j * c.kc, c.k * 8, c.h + j * c.N * c.kc, c.kc * 8, c.kc * 8, c.N, cudaMemcpyHostToDevice, c.st);
        break;
    case 3:   // D2H, device pitched columns -> host pitched columns
        for (size_t j = 0; j < nb; ++j)
            cudaMemcpy2DAsync(c.h + j * c.kc, c.k * 8, c.d + j * c.kc, c.k * 8, c.kc * 8, c.N, cudaMemcpyDeviceToHost, c.st);
        break;
    case 4:   // D2H, device pitched columns -> host dense blocks
        for (size_t j = 0; j < nb; ++j)
            cudaMemcpy2DAsync(c.h + j * c.N * c.kc, c.kc * 8, c.d + j * c.kc, c.k * 8, c.kc * 8, c.N, cudaMemcpyDeviceToHost, c.st);
        break;
    case 5:   // H2D, host pitched columns -> device pitched columns
        for (size_t j = 0; j < nb; ++j)
            cudaMemcpy2DAsync(c.d + j * c.kc, c.k * 8, c.h + j * c.kc, c.k * 8, c.kc * 8, c.N, cudaMemcpyHostToDevice, c.st);
        break;
    }
}

int main()
{
    Ctx c;
    c.N = 59711; c.k = 256;
    cudaMallocHost(&c.h, c.N * c.k * 8);
    cudaMalloc(&c.d, c.N * c.k * 8);
    cudaStreamCreate(&c.st);
    const char *names[] = {"H2D contiguous", "D2H contiguous", "H2D host dense blocks -> device pitched",
                           "D2H device pitched -> host pitched", "D2H device pitched -> host dense blocks",
                           "H2D host pitched -> device pitched"};
    for (int mode = 0; mode < 6; ++mode)
        for (size_t kc : {(size_t)64, (size_t)128}) {
            if (mode < 2 && kc != 64) continue;
            c.mode = mode; c.kc = kc;
            float ms = timeit(c.st, 5, run, &c);
            printf("{\"copy\": \"%s\", \"block_cols\": %zu, \"ms\": %.3f, \"GBps\": %.1f}\n", names[mode], kc, ms, c.N * c.k * 8 / (ms * 1e-3) / 1e9);
        }
    return 0;
}
