// K1: Newton-basis evaluation, tree (nested) form.
//   replaces eval_newton_polynomials, src/minterpy/utils.py:247-343, whose inner loops are
//   src/minterpy/jit_compiled_utils.py:62-126.
//
// p(x) = sum_{e_{m-1}} T_{m-1}[e_{m-1}] ( ... sum_{e_1} T_1[e_1] ( sum_{e_0} c_e T_0[e_0] ) )
// with T_i[e] = prod_{t<e} (x_i - gp[t,i]) -- the same chained products the reference keeps
// in `products_placeholder`.  One thread owns R query points:
//   * T_0 (dimension-0 table) lives in registers, statically indexed: every leaf (a run of
//     monomials that differ only in e_0) is one FMA per monomial, entered through a switch
//     on the leaf length;
//   * T_1 lives in shared memory as [e][r][thread] (conflict free), one FMA per leaf;
//   * dimensions >= 2 are touched once per ~100 monomials: their running product, point
//     coordinate and partial sum sit in shared memory as well;
//   * the leaf records and the packed coefficients are identical for all points: a CTA
//     streams them chunk by chunk from L2 into a 3-deep shared-memory ring with
//     cp.async.bulk + mbarrier (TMA bulk copy), and every lane reads them as broadcasts.
// All control flow depends on the index set only, never on the point: no divergence.
#include "mpx_eval_kernel.cuh"


// coefficients == nullptr: the plan's scratch already holds the packed image of these k
// polynomials (mpx_eval_host packs once, then evaluates block after block)
int mpx_launch_eval(mpx_plan *pl, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                    const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                    cudaStream_t stream)
{
    if (P == 0 && coeffs == nullptr) return MPX_OK;
    if (mpx_1d_eligible(pl, k))
        return mpx_launch_eval_1d(pl, img, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, stream);
    if (mpx_mma_eligible(pl, k))
        return mpx_launch_eval_mma(pl, img, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, stream);
    if (mpx_tmma_eligible(pl, k))
        return mpx_launch_eval_tmma(pl, img, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, stream);
    if (!pl->tree_ok)
        return mpx_fail(MPX_ERR_UNSUPPORTED,
                        "dimension-0 exponent > 64 needs the shared-memory-table kernel, whose tables "
                        "(m <= 8, m * (largest exponent + 1) * 36 doubles) do not fit for this index set");
    return tree_launch<false>(pl, img, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, stream, nullptr, nullptr);
}

// Test hook (not in the public header): ring depth of the tree kernel, 0 = the default.  The
// ring stress test runs many-chunk programs through 2-, 3- and 4-deep rings of the one-CTA
// variant.
int mpx_test_eval_stages = 0;
extern "C" void mpx_debug_set_eval_stages(int stages) { mpx_test_eval_stages = stages; }
int mpx_test_eval_uniform = 0;
extern "C" void mpx_debug_set_eval_uniform(int mode) { mpx_test_eval_uniform = mode; }

size_t mpx_eval_scratch_bytes(const mpx_plan *pl, int k)
{
    if (mpx_1d_eligible(pl, k)) return mpx_1d_scratch_bytes(pl, k);
    if (mpx_mma_eligible(pl, k)) return mpx_mma_scratch_bytes(pl, k);
    if (mpx_tmma_eligible(pl, k)) return mpx_tmma_scratch_bytes(pl, k);
    return (size_t)pl->prog.n_chunks * MPX_COEF_CAP * (size_t)k * sizeof(double);
}
