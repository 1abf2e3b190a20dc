// FP64 peak microbenchmarks for the roofline denominators (not in MEASURED_PEAKS.json):
// DFMA (vector pipe), DMMA (mma.sync f64 tensor path) and both interleaved.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int CH>
__global__ void dfma_kernel(double *out, int iters, double a, double b)
{
    double acc[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c] = threadIdx.x * 1e-9 + c;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) acc[c] = fma(acc[c], a, b);
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += acc[c];
    if (s == 123.456) out[0] = s;
}

// DFMA fed by shared-memory broadcast operands, like the evaluation kernel's leaf loop:
// one LDS.128 per 2*R FMAs
template <int R>
__global__ void dfma_lds_kernel(double *out, int iters, double a)
{
    __shared__ double2 cs[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) cs[i] = make_double2(1.0 + i * 1e-9, 0.5);
    __syncthreads();
    double acc[R], t0[R], t1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { acc[r] = r; t0[r] = a + r; t1[r] = a - r; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            double2 c = cs[(i * 16 + u) & 511];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc[r] = fma(c.y, t1[r], acc[r]);
                acc[r] = fma(c.x, t0[r], acc[r]);
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) s += acc[r];
    if (s == 123.456) out[0] = s;
}

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16816(double (&d)[4], const double (&a)[8], const double (&b)[4])
{
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, "
        "{%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
        : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
        : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
          "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int CH>
__global__ void dmma884_kernel(double *out, int iters, double a, double b)
{
    double d0[CH], d1[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) { d0[c] = c; d1[c] = -c; }
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) dmma884(d0[c], d1[c], a, b);
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += d0[c] + d1[c];
    if (s == 123.456) out[0] = s;
}

template <int CH>
__global__ void dmma16816_kernel(double *out, int iters, double av, double bv)
{
    double d[CH][4], a[8], b[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = av + i;
#pragma unroll
    for (int i = 0; i < 4; ++i) b[i] = bv - i;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
        for (int i = 0; i < 4; ++i) d[c][i] = c + i;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) dmma16816(d[c], a, b);
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += d[c][0] + d[c][1] + d[c][2] + d[c][3];
    if (s == 123.456) out[0] = s;
}

// interleaved: per iteration CH DMMA m16n8k16 and CF*... DFMAs on separate accumulators
template <int CH, int CF>
__global__ void mixed_kernel(double *out, int iters, double av, double bv)
{
    double d[CH][4], a[8], b[4], f[CF];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = av + i;
#pragma unroll
    for (int i = 0; i < 4; ++i) b[i] = bv - i;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
        for (int i = 0; i < 4; ++i) d[c][i] = c + i;
#pragma unroll
    for (int c = 0; c < CF; ++c) f[c] = c + threadIdx.x;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int c = 0; c < CH; ++c) dmma16816(d[c], a, b);
#pragma unroll
            for (int c = 0; c < CF; ++c) f[c] = fma(f[c], av, bv);
        }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += d[c][0] + d[c][1] + d[c][2] + d[c][3];
#pragma unroll
    for (int c = 0; c < CF; ++c) s += f[c];
    if (s == 123.456) out[0] = s;
}

template <class F>
double time_ms(F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    double *out;
    cudaMalloc(&out, 64);
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    const int iters = 4096;
    printf("{\"device\": \"%s\", \"sms\": %d}\n", prop.name, sms);
#define RUN(label, kern, blocks_per_sm, threads, flops_per_thread_iter)                         \
    {                                                                                          \
        double ms = time_ms([&] { kern<<<sms * (blocks_per_sm), threads>>>(out, iters, 1.0000001, 1e-9); }); \
        double fl = (double)sms * (blocks_per_sm) * (threads) * (double)iters * (flops_per_thread_iter); \
        printf("{\"bench\": \"%s\", \"blocks_per_sm\": %d, \"threads\": %d, \"ms\": %.4f, \"tflops\": %.3f}\n", \
               label, blocks_per_sm, threads, ms, fl / ms * 1e-9);                              \
    }
    RUN("dfma_ch1", dfma_kernel<1>, 2, 1024, 8 * 1 * 2.0)
    RUN("dfma_ch2", dfma_kernel<2>, 2, 1024, 8 * 2 * 2.0)
    RUN("dfma_ch4", dfma_kernel<4>, 2, 1024, 8 * 4 * 2.0)
    RUN("dfma_ch8", dfma_kernel<8>, 2, 512, 8 * 8 * 2.0)
    RUN("dfma_ch4_w8", dfma_kernel<4>, 1, 256, 8 * 4 * 2.0)
    RUN("dfma_ch4_w16", dfma_kernel<4>, 2, 256, 8 * 4 * 2.0)
    RUN("dfma_ch2_w16", dfma_kernel<2>, 2, 256, 8 * 2 * 2.0)
    RUN("dfma_ch2_w8", dfma_kernel<2>, 1, 256, 8 * 2 * 2.0)
    RUN("dfma_ch1_w16", dfma_kernel<1>, 2, 256, 8 * 1 * 2.0)
    RUN("dfma_ch1_w4", dfma_kernel<1>, 1, 128, 8 * 1 * 2.0)
#undef RUN
#define RUN1(label, kern, blocks_per_sm, threads, flops_per_thread_iter)                        \
    {                                                                                          \
        double ms = time_ms([&] { kern<<<sms * (blocks_per_sm), threads>>>(out, iters, 1.0000001); }); \
        double fl = (double)sms * (blocks_per_sm) * (threads) * (double)iters * (flops_per_thread_iter); \
        printf("{\"bench\": \"%s\", \"blocks_per_sm\": %d, \"threads\": %d, \"ms\": %.4f, \"tflops\": %.3f}\n", \
               label, blocks_per_sm, threads, ms, fl / ms * 1e-9);                              \
    }
    RUN1("dfma_lds_r1_w16", dfma_lds_kernel<1>, 2, 256, 16 * 2 * 1 * 2.0)
    RUN1("dfma_lds_r2_w16", dfma_lds_kernel<2>, 2, 256, 16 * 2 * 2 * 2.0)
    RUN1("dfma_lds_r4_w8", dfma_lds_kernel<4>, 1, 256, 16 * 2 * 4 * 2.0)
    RUN1("dfma_lds_r2_w8", dfma_lds_kernel<2>, 1, 256, 16 * 2 * 2 * 2.0)
#undef RUN1
#define RUN(label, kern, blocks_per_sm, threads, flops_per_warp_iter)                           \
    {                                                                                          \
        double ms = time_ms([&] { kern<<<sms * (blocks_per_sm), threads>>>(out, iters, 1.0000001, 1e-9); }); \
        double fl = (double)sms * (blocks_per_sm) * ((threads) / 32) * (double)iters * (flops_per_warp_iter); \
        printf("{\"bench\": \"%s\", \"blocks_per_sm\": %d, \"threads\": %d, \"ms\": %.4f, \"tflops\": %.3f}\n", \
               label, blocks_per_sm, threads, ms, fl / ms * 1e-9);                              \
    }
    RUN("dmma884_ch4", dmma884_kernel<4>, 2, 512, 8 * 4 * (2.0 * 8 * 8 * 4))
    RUN("dmma884_ch8", dmma884_kernel<8>, 2, 512, 8 * 8 * (2.0 * 8 * 8 * 4))
    RUN("dmma16816_ch2", dmma16816_kernel<2>, 2, 512, 4 * 2 * (2.0 * 16 * 8 * 16))
    RUN("dmma16816_ch4", dmma16816_kernel<4>, 2, 512, 4 * 4 * (2.0 * 16 * 8 * 16))
    RUN("dmma16816_ch4_w8", dmma16816_kernel<4>, 1, 256, 4 * 4 * (2.0 * 16 * 8 * 16))
    // mixed: flops = DMMA + DFMA parts
    RUN("mixed_mma4_fma4", (mixed_kernel<4, 4>), 2, 512, 4 * (4 * (2.0 * 16 * 8 * 16) + 4 * 32 * 2.0))
    RUN("mixed_mma4_fma16", (mixed_kernel<4, 16>), 2, 512, 4 * (4 * (2.0 * 16 * 8 * 16) + 16 * 32 * 2.0))
    RUN("mixed_mma2_fma32", (mixed_kernel<2, 32>), 2, 512, 4 * (2 * (2.0 * 16 * 8 * 16) + 32 * 32 * 2.0))
#undef RUN
    cudaFree(out);
    return 0;
}
