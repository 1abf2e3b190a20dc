// K1m: the (P x N) Newton-monomial matrix.
//   replaces eval_newton_monomials, src/minterpy/utils.py:177-244, i.e.
//   eval_newton_monomials_multiple / _single, src/minterpy/jit_compiled_utils.py:62-211.
// Write-bandwidth bound (8 B per point and monomial).  A CTA owns PT = 8 points: their product
// tables go to shared memory as [dimension][exponent][point], then consecutive threads take
// consecutive monomials (the stores of one output row are coalesced) and each thread forms
// its monomial for all 8 points: one 8-byte load of the packed exponents, four 16-byte table
// reads per dimension, 8 multiplications per further dimension, 8 stores.
// Rows of the (P x N) result start wherever N puts them (N is usually odd), and warp stores
// that straddle 128-byte lines cost a third of the bandwidth (measured: 3.7 vs 5.6 TB/s at
// N = 24809).  So a CTA takes 8 points that are 16 rows apart -- their rows start at the same
// offset modulo 128 bytes -- and shifts its monomial range left by that offset: every warp
// store then covers whole lines, in all 8 rows.  Same multiplication
// order as the reference (dimension 0 first, no FMA possible) => bit-identical values; a
// factor T[0] = 1.0 is an exact multiplication, so skipping zero exponents (reference :124)
// needs no branch, and 1.0 * T0 is T0.
#include "mpx_common.h"

namespace {

constexpr int PT = 8;    // points per CTA
constexpr int TP = 10;   // doubles per (dimension, exponent) row: 8 points + 2 of padding, so
                         // that 8 consecutive exponents fall into 8 different bank groups

template <int M, bool TRI>
__global__ void __launch_bounds__(256)
monomials_kernel(const double *__restrict__ xx, int64_t ld_xx, int64_t P,
                 const uint64_t *__restrict__ exp64, int64_t N,
                 const double *__restrict__ gp, int gp_rows, const int *__restrict__ max_exp_d,
                 double *__restrict__ out, int64_t ld_out, int64_t n_per_cta)
{
    extern __shared__ __align__(16) double tab[];   // [M][gp_rows + 1][TP]
    const int TR = gp_rows + 1;
    // points p0, p0 + 16, .., p0 + 112 of a block of 128 consecutive points
    const int64_t p0 = ((int64_t)(blockIdx.x >> 4) * (16 * PT)) + (blockIdx.x & 15);
    if (p0 >= P) return;
    const int n_valid = (P - p0 >= 16 * PT) ? PT : (int)((P - p0 + 15) / 16);
    for (int t = threadIdx.x; t < PT * M; t += blockDim.x) {
        const int pt = t % PT, i = t / PT;
        double *T = tab + (size_t)i * TR * TP + pt;
        T[0] = 1.0;
        const int64_t p = (pt < n_valid) ? p0 + 16 * pt : p0;  // stand-in for the ragged tail
        const double x = xx[p * ld_xx + i];
        double prod = 1.0;
        const int me = max_exp_d[i];
        for (int j = 0; j < me; ++j) {           // jit_compiled_utils.py:102-112
            prod = __dmul_rn(prod, __dsub_rn(x, gp[(size_t)i * gp_rows + j]));
            T[(size_t)(j + 1) * TP] = prod;
        }
    }
    __syncthreads();
    const int64_t j0 = (int64_t)blockIdx.y * n_per_cta;
    const int64_t j1 = (j0 + n_per_cta < N) ? j0 + n_per_cta : N;
    double *orow = out + p0 * ld_out;
    const int64_t row_stride = 16 * ld_out;
    // first monomial of thread 0: the last one at or before j0 whose address is 128-byte aligned
    const int64_t a = (int64_t)((reinterpret_cast<uintptr_t>(orow) >> 3) & 15);
    int64_t j = j0 - ((a + j0) & 15) + threadIdx.x;
    if (j >= j1) return;
    uint64_t e = exp64[j < j0 ? j0 : j];
    for (; j < j1; j += 256) {
        const int64_t jn = j + 256;
        const uint64_t e_next = exp64[jn < j1 ? jn : j1 - 1];   // next round's exponents
        double v[PT];
        {
            const double2 *r = reinterpret_cast<const double2 *>(tab + (size_t)(e & 255u) * TP);
#pragma unroll
            for (int h = 0; h < PT / 2; ++h) {
                const double2 a = r[h];
                v[2 * h] = a.x;
                v[2 * h + 1] = a.y;
            }
        }
#pragma unroll
        for (int i = 1; i < M; ++i) {            // :115-126
            const uint32_t ei = (uint32_t)(e >> (8 * i)) & 255u;
            const double2 *r = reinterpret_cast<const double2 *>(tab + ((size_t)i * TR + ei) * TP);
#pragma unroll
            for (int h = 0; h < PT / 2; ++h) {
                const double2 a = r[h];
                v[2 * h] = __dmul_rn(v[2 * h], a.x);
                v[2 * h + 1] = __dmul_rn(v[2 * h + 1], a.y);
            }
        }
#pragma unroll
        for (int pt = 0; pt < PT; ++pt) {
            if (TRI && j > p0 + 16 * pt) v[pt] = 0.0;            // :187-199
            if (pt < n_valid && j >= j0) __stcs(orow + pt * row_stride + j, v[pt]);
        }
        e = e_next;
    }
}

// any number of dimensions (m > 8): exponents as bytes, one table read per point and dimension
__global__ void __launch_bounds__(256)
monomials_generic_kernel(const double *__restrict__ xx, int64_t ld_xx, int64_t P,
                         const uint8_t *__restrict__ exp8, int64_t N, int m,
                         const double *__restrict__ gp, int gp_rows,
                         const int *__restrict__ max_exp_d, double *__restrict__ out,
                         int64_t ld_out, int triangular, int64_t n_per_cta)
{
    extern __shared__ __align__(16) double tab[];   // [PT][m][gp_rows + 1]
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
            for (int j = 0; j < me; ++j) {
                prod = __dmul_rn(prod, __dsub_rn(x, gp[(size_t)i * gp_rows + j]));
                T[j + 1] = prod;
            }
        }
    }
    __syncthreads();
    const int64_t j0 = (int64_t)blockIdx.y * n_per_cta;
    const int64_t j1 = (j0 + n_per_cta < N) ? j0 + n_per_cta : N;
    for (int64_t j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
        for (int pt = 0; pt < PT; ++pt) {
            const int64_t p = p0 + pt;
            if (p >= P) break;
            const double *T = tab + (size_t)pt * m * TR;
            double v = 1.0;
            for (int i = 0; i < m; ++i) v = __dmul_rn(v, T[i * TR + exp8[j * m + i]]);
            if (triangular && j > p) v = 0.0;
            out[p * ld_out + j] = v;
        }
    }
}

template <int M>
void launch_m(bool tri, dim3 grid, size_t smem, cudaStream_t st, const double *xx, int64_t ld_xx,
              int64_t P, const mpx_plan *pl, double *out, int64_t ld_out, int64_t n_per_cta)
{
    if (tri) {
        mpx_allow_max_smem(monomials_kernel<M, true>);
        monomials_kernel<M, true><<<grid, 256, smem, st>>>(xx, ld_xx, P, pl->d_exp64, pl->N, pl->d_gp,
                                                          pl->gp_rows, pl->d_max_exp, out, ld_out, n_per_cta);
    } else {
        mpx_allow_max_smem(monomials_kernel<M, false>);
        monomials_kernel<M, false><<<grid, 256, smem, st>>>(xx, ld_xx, P, pl->d_exp64, pl->N, pl->d_gp,
                                                           pl->gp_rows, pl->d_max_exp, out, ld_out, n_per_cta);
    }
}

}  // namespace

int mpx_launch_monomials(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx, double *out,
                         int64_t ld_out, int triangular, cudaStream_t stream)
{
    if (P == 0) return MPX_OK;
    const bool packed = pl->m <= 8 && pl->d_exp64 != nullptr;
    const size_t smem = packed ? (size_t)pl->m * (pl->gp_rows + 1) * TP * sizeof(double)
                               : (size_t)PT * pl->m * (pl->gp_rows + 1) * sizeof(double);
    if (smem > 200 * 1024) return mpx_fail(MPX_ERR_UNSUPPORTED, "tables too large for shared memory");
    // packed kernel: 16 CTAs per block of 128 points (see the kernel); generic: 8 points per CTA
    const int64_t ptiles = packed ? (P + 16 * PT - 1) / (16 * PT) * 16 : (P + PT - 1) / PT;
    if (ptiles > 0x7fffffffLL) return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points for one launch");
    // split the monomials over grid.y only when there are too few point tiles to fill the GPU
    int64_t ysplit = 1;
    if (ptiles < 2 * 148) ysplit = std::min<int64_t>((2 * 148 + ptiles - 1) / ptiles, (pl->N + 255) / 256);
    if (ysplit < 1) ysplit = 1;
    const int64_t n_per_cta = (pl->N + ysplit - 1) / ysplit;
    dim3 grid((unsigned)ptiles, (unsigned)ysplit, 1);
    if (packed) {
        const bool tri = triangular != 0;
        switch (pl->m) {
        case 1: launch_m<1>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        case 2: launch_m<2>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        case 3: launch_m<3>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        case 4: launch_m<4>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        case 5: launch_m<5>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        case 6: launch_m<6>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        case 7: launch_m<7>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        default: launch_m<8>(tri, grid, smem, stream, xx, ld_xx, P, pl, out, ld_out, n_per_cta); break;
        }
    } else {
        MPX_CUDA(mpx_allow_max_smem(monomials_generic_kernel));
        monomials_generic_kernel<<<grid, 256, smem, stream>>>(xx, ld_xx, P, pl->d_exp8, pl->N, pl->m,
                                                             pl->d_gp, pl->gp_rows, pl->d_max_exp,
                                                             out, ld_out, triangular, n_per_cta);
    }
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}
