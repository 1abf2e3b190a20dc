// Probe: how much FP64-pipe throughput does a shared-memory load cost when it is interleaved
// with DFMA chains?  (Design evidence for the evaluation kernel's leaf loop; not product code.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/lds_dfma_probe tools/lds_dfma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double2 lds128(unsigned a)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ double lds64(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}

__constant__ double2 g_cst[4096];     // 64 KB: the whole user constant bank

// MODE 7: operands from the constant bank at a warp-uniform running index (streams through 64 KB)
// MODE 0: no load (operands in registers); 1: one broadcast LDS.128 per 2R DFMA; 2: two broadcast
// LDS.64 per 2R DFMA; 3: one per-lane LDS.128 per 2R DFMA; 4: one broadcast LDS.64 per 2R DFMA
// (half the bytes); 5: integer ALU op instead of the load; 6: LDS.128 per 4R DFMA
template <int R, int MODE>
__global__ void probe(double *out, int iters, double a)
{
    __shared__ double2 cs[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) cs[i] = make_double2(1.0 + i * 1e-9, 0.5);
    __syncthreads();
    double acc[R], t0[R], t1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { acc[r] = r; t0[r] = a + r; t1[r] = a - r; }
    const unsigned base = (unsigned)__cvta_generic_to_shared(cs);
    const unsigned lane_off = (MODE == 3) ? (threadIdx.x & 31) * 16 : 0;
    unsigned junk = threadIdx.x;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            double2 c = make_double2(a, a);
            const unsigned ad = base + lane_off + (((i * 16 + u) & 31) * 512);
            if (MODE == 1 || MODE == 3) c = lds128(ad);
            if (MODE == 2) { c.x = lds64(ad); c.y = lds64(ad + 8); }
            if (MODE == 4) { c.x = lds64(ad); c.y = c.x; }
            if (MODE == 5) { junk = junk * 3 + u; }
            if (MODE == 7) { c = g_cst[(i * 16 + u + blockIdx.x * 7) & 4095]; }
            if (MODE == 8) { c = g_cst[(i * 16 + u) & 255]; }
            if (MODE == 6) { if ((u & 1) == 0) c = lds128(ad); else c = make_double2(acc[0] * 0 + a, a); }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc[r] = fma(c.y, t1[r], acc[r]);
                acc[r] = fma(c.x, t0[r], acc[r]);
            }
        }
    }
    double s = junk;
#pragma unroll
    for (int r = 0; r < R; ++r) s += acc[r];
    if (s == 123.456) out[0] = s;
}

template <class F>
void run(const char *name, F kernel, int threads, int blocks_per_sm, int iters, double flops_per_thread_iter)
{
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double *out;
    cudaMalloc(&out, 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sms * blocks_per_sm;
    kernel<<<grid, threads>>>(out, iters, 1.000001);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        kernel<<<grid, threads>>>(out, iters, 1.000001);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double fl = flops_per_thread_iter * iters * (double)threads * grid;
    printf("{\"bench\": \"%s\", \"threads\": %d, \"blocks_per_sm\": %d, \"ms\": %.4f, \"tflops\": %.3f}\n", name,
           threads, blocks_per_sm, best, fl / (best * 1e-3) / 1e12);
    cudaFree(out);
}

int main()
{
    const int it = 20000;
    {
        static double2 h[4096];
        for (int i = 0; i < 4096; ++i) h[i] = make_double2(1.0 + i * 1e-9, 0.5);
        cudaMemcpyToSymbol(g_cst, h, sizeof(h));
    }
    // 8 warps per SM (2 per scheduler), like the evaluation kernel at cfg2
    run("r4_none_w8", probe<4, 0>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_lds128_bcast_w8", probe<4, 1>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_2xlds64_bcast_w8", probe<4, 2>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_lds128_lane_w8", probe<4, 3>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_lds64_bcast_w8", probe<4, 4>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_ialu_w8", probe<4, 5>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_lds128_per16_w8", probe<4, 6>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_const64k_w8", probe<4, 7>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_const4k_w8", probe<4, 8>, 256, 1, it, 16.0 * 4 * 4);
    run("r4_const64k_w16", probe<4, 7>, 256, 2, it, 16.0 * 4 * 4);
    // 16 warps per SM
    run("r4_none_w16", probe<4, 0>, 256, 2, it, 16.0 * 4 * 4);
    run("r4_lds128_bcast_w16", probe<4, 1>, 256, 2, it, 16.0 * 4 * 4);
    run("r2_none_w16", probe<2, 0>, 256, 2, it, 16.0 * 4 * 2);
    run("r2_lds128_bcast_w16", probe<2, 1>, 256, 2, it, 16.0 * 4 * 2);
    run("r2_lds128_bcast_w32", probe<2, 1>, 512, 2, it, 16.0 * 4 * 2);
    run("r8_none_w8", probe<8, 0>, 256, 1, it, 16.0 * 4 * 8);
    run("r8_lds128_bcast_w8", probe<8, 1>, 256, 1, it, 16.0 * 4 * 8);
    run("r4_lds128_bcast_w4", probe<4, 1>, 128, 1, it, 16.0 * 4 * 4);
    run("r4_none_w4", probe<4, 0>, 128, 1, it, 16.0 * 4 * 4);
    return 0;
}
