// Partial derivatives of few polynomials on the tree kernel (K1).
//   replaces deriv_newt_eval, src/minterpy/polynomials/utils.py:8-112 (the evaluation behind
//   NewtonPolynomial.diff / partial_diff, src/minterpy/polynomials/newton_polynomial.py:55-86).
//
// A partial derivative only changes the per-dimension tables: entry q of dimension i becomes
// d^o/dx^o prod_{t<q}(x_i - g_t).  The tree kernel keeps dimension 0 in registers and dimension 1
// in shared memory -- both are rewritten in place by the product-rule recurrence when the kernel
// starts (DERIV instantiation, mpx_eval_kernel.cuh).  Dimensions >= 2 it normally follows with a
// RUNNING product, which a derivative breaks; for those a small kernel here tabulates every entry
// for every point of the launch in global memory (coalesced by point), and the folds of the DERIV
// kernel look the entry up instead of multiplying: one 8-byte load per point and record.
// Many polynomials (k >= MPX_INFO_MMA_MIN_K) stay on the tensor-core kernel (mpx_eval_mma.cu).
#include "mpx_eval_kernel.cuh"

namespace {

// table entry e of dimension 2+u for point p at tab[(u * rows + e) * P + p]
__global__ void deriv_tables_kernel(const double *__restrict__ xx, int64_t ld_xx, int64_t P,
                                    const double *__restrict__ gp, int gp_rows, int m,
                                    EvalParams prm, double *__restrict__ tab)
{
    const int64_t pt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pt >= P) return;
    for (int u = 0; u + 2 < m; ++u) {
        const int dim = 2 + u, me = prm.max_exp[dim], order = prm.order[dim];
        const double x = xx[pt * ld_xx + dim];
        const double *g = gp + (size_t)dim * gp_rows;
        double *T = tab + (size_t)u * gp_rows * P + pt;              // T[e * P]
        double prod = 1.0;
        T[0] = 1.0;
        for (int j = 0; j < me; ++j) {
            prod *= x - g[j];
            T[(size_t)(j + 1) * P] = prod;
        }
        // F_j[q+1] = (x - g_q) F_j[q] + F_{j-1}[q], one pass per order, in place; then o!
        double fact = 1.0;
        for (int pass = 1; pass <= order; ++pass) {
            double fj = 0.0, old = T[0];
            T[0] = 0.0;
            for (int q = 0; q < me; ++q) {
                const double old_next = T[(size_t)(q + 1) * P];
                fj = __fma_rn(x - g[q], fj, old);
                T[(size_t)(q + 1) * P] = fj;
                old = old_next;
            }
            fact *= (double)pass;
        }
        if (order > 0)
            for (int q = 0; q <= me; ++q) T[(size_t)q * P] *= fact;
    }
}

}  // namespace

int mpx_launch_eval_tree_deriv(mpx_plan *pl, MpxImage *img, MpxImage *tables, const double *xx,
                               int64_t P, int64_t ld_xx, const double *coeffs, int k, int64_t ld_c,
                               const int *orders_h, double *out, int64_t ld_out, cudaStream_t stream)
{
    const int U = pl->m > 2 ? pl->m - 2 : 0;
    // points per launch: bounds the table scratch (U * gp_rows * 8 MiB)
    const int64_t chunk = (int64_t)1 << 20;
    EvalParams tp;
    for (int i = 0; i < MPX_MAX_M; ++i) {
        tp.max_exp[i] = pl->max_exp[i];
        tp.order[i] = i < pl->m ? orders_h[i] : 0;
    }
    for (int64_t p0 = 0; p0 < P; p0 += chunk) {
        const int64_t np = std::min(chunk, P - p0);
        const double *x0 = xx + p0 * ld_xx;
        double *tab = nullptr;
        if (U > 0) {
            int rc = mpx_image_acquire(tables, (size_t)U * pl->gp_rows * np * sizeof(double), stream);
            if (rc != MPX_OK) return rc;
            tab = tables->ptr;
            deriv_tables_kernel<<<(unsigned)((np + 255) / 256), 256, 0, stream>>>(
                x0, ld_xx, np, pl->d_gp, pl->gp_rows, pl->m, tp, tab);
            MPX_CUDA(cudaGetLastError());
        }
        int rc = tree_launch<true>(pl, img, x0, np, ld_xx, p0 == 0 ? coeffs : nullptr, k, ld_c,
                                   out + p0 * ld_out, ld_out, stream, orders_h, tab);
        if (rc != MPX_OK) return rc;
        if (U > 0 && (rc = mpx_image_release(tables, stream)) != MPX_OK) return rc;
    }
    return MPX_OK;
}
