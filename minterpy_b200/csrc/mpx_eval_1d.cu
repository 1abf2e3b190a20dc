// K1h: one-dimensional Newton polynomials by Horner's rule.
//   replaces eval_newton_polynomials, src/minterpy/utils.py:247-343, for m = 1 and any number of
//   polynomials -- the README quickstart and BASELINE config 1 (degree 64, 1e6 points).
//
// p(x) = c_0 + (x - g_0) (c_1 + (x - g_1) (c_2 + ... + (x - g_{n-1}) c_n)): the same monomials
// prod_{t<e} (x - g_t) the reference tabulates in `products_placeholder`
// (jit_compiled_utils.py:102-112, rows t < max exponent), summed from the top.  One subtraction
// and one FMA per coefficient, no table: a thread keeps only its points and running sums, so
// the SM is full of warps -- the tree kernel (mpx_eval.cu) needs 2 x 65 registers for the
// dimension-0 table of a degree-64 polynomial, runs 8 warps per SM and was bound by the latency
// of its 64-step table chain at 0.1 of the FP64 pipe.
// Exponents may be unsorted and have gaps: the pack kernel scatters the coefficients into a
// dense (max exponent + 1) x k image, zero where an exponent is absent.
// Values agree with the tree kernel within rounding (different summation order), like every
// evaluation kernel here against the reference's BLAS product.
#include "mpx_common.h"

int mpx_test_eval_1d = 0;      // test hook: 2 = never (m = 1 runs on the tree kernel)
extern "C" void mpx_debug_set_eval_1d(int mode) { mpx_test_eval_1d = mode; }

namespace {

constexpr int H1_THREADS = 256;
constexpr int H1_MAX_K = 4 * 65535;   // polynomials go four to a grid row

__global__ void pack_dense_1d_kernel(const double *__restrict__ coeffs, int64_t ld_c, int k,
                                     const int32_t *__restrict__ row_of_exp, int n1,
                                     double *__restrict__ dense)
{
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= (int64_t)n1 * k) return;
    const int j = (int)(i / k), col = (int)(i % k);
    const int32_t r = row_of_exp[j];
    dense[i] = r >= 0 ? coeffs[(int64_t)r * ld_c + col] : 0.0;
}

// R points per thread, KB polynomials per CTA row (blockIdx.y).  Shared memory holds one record
// per exponent j: KB coefficients, then g_j, padded to an even number of doubles -- KB = 1: one
// 16-byte broadcast load per step for 2 R FP64 instructions.
template <int R, int KB>
__global__ void __launch_bounds__(H1_THREADS)
horner1d_kernel(const double *__restrict__ xx, int64_t ld_xx, int64_t P,
                const double *__restrict__ dense, int k, const double *__restrict__ g, int n1,
                double *__restrict__ out, int64_t ld_out)
{
    constexpr int REC = (KB + 2) & ~1;
    extern __shared__ __align__(16) double rec[];
    const int col0 = blockIdx.y * KB;
    for (int i = threadIdx.x; i < n1 * REC; i += H1_THREADS) {
        const int j = i / REC, c = i % REC;
        double v = 0.0;
        if (c < KB) v = (col0 + c < k) ? dense[(int64_t)j * k + col0 + c] : 0.0;
        else if (c == KB && j + 1 < n1) v = g[j];       // rows < max exponent are read, as upstream
        rec[i] = v;
    }
    const int64_t base = (int64_t)blockIdx.x * (H1_THREADS * R) + threadIdx.x;
    double x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t p = base + (int64_t)r * H1_THREADS;
        x[r] = p < P ? xx[p * ld_xx] : 0.0;
    }
    __syncthreads();
    double acc[R][KB];
    {
        const double *top = rec + (size_t)(n1 - 1) * REC;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int c = 0; c < KB; ++c) acc[r][c] = top[c];
    }
#pragma unroll 4
    for (int j = n1 - 2; j >= 0; --j) {
        const double2 *rj = reinterpret_cast<const double2 *>(rec + (size_t)j * REC);
        double cj[REC];
#pragma unroll
        for (int q = 0; q < REC / 2; ++q) {
            const double2 t = rj[q];
            cj[2 * q] = t.x;
            cj[2 * q + 1] = t.y;
        }
        const double gj = cj[KB];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double d = __dsub_rn(x[r], gj);
#pragma unroll
            for (int c = 0; c < KB; ++c) acc[r][c] = __fma_rn(acc[r][c], d, cj[c]);
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t p = base + (int64_t)r * H1_THREADS;
        if (p < P) {
#pragma unroll
            for (int c = 0; c < KB; ++c)
                if (col0 + c < k) out[p * ld_out + col0 + c] = acc[r][c];
        }
    }
}

template <int R, int KB>
int launch_horner(const mpx_plan *pl, const double *dense, const double *xx, int64_t P, int64_t ld_xx,
                  int k, double *out, int64_t ld_out, cudaStream_t stream)
{
    const int n1 = pl->max_exp[0] + 1;
    constexpr int REC = (KB + 2) & ~1;
    const int64_t tiles = (P + (int64_t)H1_THREADS * R - 1) / ((int64_t)H1_THREADS * R);
    if (tiles > 0x7fffffffLL) return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points for one launch");
    dim3 grid((unsigned)tiles, (unsigned)((k + KB - 1) / KB), 1);
    horner1d_kernel<R, KB><<<grid, H1_THREADS, (size_t)n1 * REC * sizeof(double), stream>>>(
        xx, ld_xx, P, dense, k, pl->d_gp, n1, out, ld_out);
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}

}  // namespace

bool mpx_1d_eligible(const mpx_plan *pl, int k)
{
    return pl->m == 1 && pl->d_row1d != nullptr && k <= H1_MAX_K && mpx_test_eval_1d != 2;
}

size_t mpx_1d_scratch_bytes(const mpx_plan *pl, int k)
{
    return (size_t)(pl->max_exp[0] + 1) * (size_t)k * sizeof(double);
}

// same protocol as the other launchers: coeffs == nullptr -> `img` already holds these k
// polynomials; P == 0 with coefficients -> pack only
int mpx_launch_eval_1d(mpx_plan *pl, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                       const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                       cudaStream_t stream)
{
    const int n1 = pl->max_exp[0] + 1;
    if (coeffs) {
        int rc = mpx_image_acquire(img, mpx_1d_scratch_bytes(pl, k), stream);
        if (rc != MPX_OK) return rc;
        const int64_t total = (int64_t)n1 * k;
        pack_dense_1d_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
            coeffs, ld_c, k, pl->d_row1d, n1, img->ptr);
        MPX_CUDA(cudaGetLastError());
    }
    if (P == 0) return coeffs ? mpx_image_release(img, stream) : MPX_OK;
    int rc;
    // (eight points per thread amortise the broadcast load; fewer when the grid would not cover
    // the SMs otherwise)
    if (k == 1 && P >= (int64_t)1 << 19) rc = launch_horner<8, 1>(pl, img->ptr, xx, P, ld_xx, k, out, ld_out, stream);
    else if (k == 1) rc = launch_horner<2, 1>(pl, img->ptr, xx, P, ld_xx, k, out, ld_out, stream);
    else if (k == 2) rc = launch_horner<4, 2>(pl, img->ptr, xx, P, ld_xx, k, out, ld_out, stream);
    else rc = launch_horner<4, 4>(pl, img->ptr, xx, P, ld_xx, k, out, ld_out, stream);
    if (rc != MPX_OK) return rc;
    return mpx_image_release(img, stream);
}
