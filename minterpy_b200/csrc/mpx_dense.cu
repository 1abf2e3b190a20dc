// K4: dense operator apply  out = T @ c   (FP64).
//   replaces MatrixOperator.__matmul__, src/minterpy/schemes/matrix_operator.py:13-22.
// T is (rows x cols); for the transformations of this path it is N x N, read once from HBM
// (8 N^2 bytes) and that read bounds the kernel for every k the library meets.  A CTA owns
// 32 rows; a warp owns 4 of them and sweeps the columns, each lane keeping KT partial sums per
// row, reduced by shuffles at the end.
#include "mpx_common.h"

namespace {

constexpr int ROWS_PER_WARP = 4;
constexpr int WARPS = 8;
constexpr int KT = 4;     // columns of c per pass
constexpr int UNR = 4;    // column strips of 64 in flight per row (double2 loads)

// lanes sweep the columns two at a time (16-byte loads), UNR strips per row are issued before
// any of them is used: 4 rows x 4 strips x 16 B = 256 B in flight per lane
__global__ void __launch_bounds__(WARPS * 32)
dense_apply_kernel(const double *__restrict__ T, int64_t rows, int64_t cols, int64_t ld_t,
                   const double *__restrict__ c, int k, int64_t ld_c, double *__restrict__ out,
                   int64_t ld_out)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t r0 = ((int64_t)blockIdx.x * WARPS + warp) * ROWS_PER_WARP;
    const int q0 = blockIdx.y * KT;
    if (r0 >= rows) return;
    double acc[ROWS_PER_WARP][KT];
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
        for (int q = 0; q < KT; ++q) acc[i][q] = 0.0;
    const bool vec_ok = ((ld_t & 1) == 0) && ((reinterpret_cast<uintptr_t>(T) & 15) == 0);
    int64_t j0 = 0;
    if (vec_ok) {
        const int64_t span = 64 * UNR;
        for (; j0 + span <= cols; j0 += span) {
            double2 t[ROWS_PER_WARP][UNR];
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                const int64_t r = (r0 + i < rows) ? r0 + i : rows - 1;
#pragma unroll
                for (int u = 0; u < UNR; ++u)
                    t[i][u] = __ldcs(reinterpret_cast<const double2 *>(T + r * ld_t + j0 + 64 * u + 2 * lane));
            }
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int64_t j = j0 + 64 * u + 2 * lane;
                double cx[KT], cy[KT];
#pragma unroll
                for (int q = 0; q < KT; ++q) {
                    cx[q] = (q0 + q < k) ? c[j * ld_c + q0 + q] : 0.0;
                    cy[q] = (q0 + q < k) ? c[(j + 1) * ld_c + q0 + q] : 0.0;
                }
#pragma unroll
                for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
                    for (int q = 0; q < KT; ++q)
                        acc[i][q] = fma(t[i][u].y, cy[q], fma(t[i][u].x, cx[q], acc[i][q]));
            }
        }
    }
    for (int64_t j = j0 + lane; j < cols; j += 32) {          // remainder / unaligned matrices
        double cv[KT];
#pragma unroll
        for (int q = 0; q < KT; ++q) cv[q] = (q0 + q < k) ? c[j * ld_c + q0 + q] : 0.0;
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i) {
            const double t = (r0 + i < rows) ? T[(r0 + i) * ld_t + j] : 0.0;
#pragma unroll
            for (int q = 0; q < KT; ++q) acc[i][q] = fma(t, cv[q], acc[i][q]);
        }
    }
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
        for (int q = 0; q < KT; ++q) {
            double v = acc[i][q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && r0 + i < rows && q0 + q < k) out[(r0 + i) * ld_out + q0 + q] = v;
        }
}

}  // namespace

extern "C" int mpx_apply_dense(int device, const double *T, int64_t rows, int64_t cols,
                               int64_t ld_t, const double *c, int k, int64_t ld_c, double *out,
                               int64_t ld_out, void *stream)
{
    if (rows < 0 || cols < 0 || k < 1) return mpx_fail(MPX_ERR_ARG, "bad sizes");
    if (rows == 0) return MPX_OK;
    if (!T || !c || !out) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld_t < cols || ld_c < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select CUDA device");
    const int64_t rows_per_cta = (int64_t)WARPS * ROWS_PER_WARP;
    dim3 grid((unsigned)((rows + rows_per_cta - 1) / rows_per_cta), (unsigned)((k + KT - 1) / KT), 1);
    dense_apply_kernel<<<grid, WARPS * 32, 0, (cudaStream_t)stream>>>(T, rows, cols, ld_t, c, k, ld_c,
                                                                    out, ld_out);
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}
