// K1m: the (P x N) Newton-monomial matrix.
//   replaces eval_newton_monomials, src/minterpy/utils.py:177-244, i.e.
//   eval_newton_monomials_multiple / _single, src/minterpy/jit_compiled_utils.py:62-211.
// Write-bandwidth bound (8 B per point and monomial).  A CTA owns PT points: their product
// tables go to shared memory, then consecutive threads take consecutive monomials so the
// stores of one output row are coalesced.  Same multiplication order as the reference
// (dimension 0 first, no FMA possible) => bit-identical values; a factor T[0] = 1.0 is an
// exact multiplication, so skipping zero exponents (reference :124) needs no branch.
#include "mpx_common.h"

namespace {

constexpr int PT = 8;   // points per CTA

__global__ void __launch_bounds__(256)
monomials_kernel(const double *__restrict__ xx, int64_t ld_xx, int64_t P,
                 const uint8_t *__restrict__ exp8, int64_t N, int m,
                 const double *__restrict__ gp, int gp_rows, const int *__restrict__ max_exp_d,
                 double *__restrict__ out, int64_t ld_out, int triangular, int64_t n_per_cta)
{
    extern __shared__ double tab[];   // [PT][m][gp_rows + 1]
    const int TR = gp_rows + 1;
    const int64_t p0 = (int64_t)blockIdx.x * PT;
    for (int t = threadIdx.x; t < PT * m; t += blockDim.x) {
        const int pt = t / m, i = t % m;
        const int64_t p = p0 + pt;
        double *T = tab + (size_t)(pt * m + i) * TR;
        T[0] = 1.0;
        if (p < P) {
            const double x = xx[p * ld_xx + i];
            double prod = 1.0;
            const int me = max_exp_d[i];
            for (int j = 0; j < me; ++j) {       // jit_compiled_utils.py:102-112
                prod = __dmul_rn(prod, __dsub_rn(x, gp[(size_t)i * gp_rows + j]));
                T[j + 1] = prod;
            }
        }
    }
    __syncthreads();
    const int64_t j0 = (int64_t)blockIdx.y * n_per_cta;
    const int64_t j1 = (j0 + n_per_cta < N) ? j0 + n_per_cta : N;
    for (int64_t j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
        int e[MPX_MAX_M];
        for (int i = 0; i < m; ++i) e[i] = exp8[j * m + i];
#pragma unroll
        for (int pt = 0; pt < PT; ++pt) {
            const int64_t p = p0 + pt;
            if (p >= P) break;
            const double *T = tab + (size_t)pt * m * TR;
            double v = 1.0;
            for (int i = 0; i < m; ++i) v = __dmul_rn(v, T[i * TR + e[i]]);   // :115-126
            if (triangular && j > p) v = 0.0;                                   // :187-199
            out[p * ld_out + j] = v;
        }
    }
}

}  // namespace

int mpx_launch_monomials(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx, double *out,
                         int64_t ld_out, int triangular, cudaStream_t stream)
{
    if (P == 0) return MPX_OK;
    const size_t smem = (size_t)PT * pl->m * (pl->gp_rows + 1) * sizeof(double);
    if (smem > 200 * 1024) return mpx_fail(MPX_ERR_UNSUPPORTED, "tables too large for shared memory");
    MPX_CUDA(cudaFuncSetAttribute(monomials_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
    const int64_t ptiles = (P + PT - 1) / PT;
    if (ptiles > 0x7fffffffLL) return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points for one launch");
    // split the monomials over grid.y only when there are too few point tiles to fill the GPU
    int64_t ysplit = 1;
    if (ptiles < 2 * 148) ysplit = std::min<int64_t>((2 * 148 + ptiles - 1) / ptiles, (pl->N + 255) / 256);
    if (ysplit < 1) ysplit = 1;
    const int64_t n_per_cta = (pl->N + ysplit - 1) / ysplit;
    dim3 grid((unsigned)ptiles, (unsigned)ysplit, 1);
    monomials_kernel<<<grid, 256, smem, stream>>>(xx, ld_xx, P, pl->d_exp8, pl->N, pl->m, pl->d_gp,
                                                 pl->gp_rows, pl->d_max_exp, out, ld_out,
                                                 triangular, n_per_cta);
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}
