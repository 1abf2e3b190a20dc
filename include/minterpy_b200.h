/*
 * minterpy_b200.h -- C ABI of libminterpy_b200.so (hand-written CUDA for sm_100a).
 *
 * The reference (casus/minterpy) is pure Python + Numba and has no FFI of its own; the
 * drop-in boundary is therefore the set of Python callables named in SURVEY.md section 8(b),
 * and this header is what a ctypes binding for those callables needs.  Every entry point
 * cites the reference function it replaces (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain C, no exceptions; every function returns an int status, 0 = MPX_OK.  The text
 *     of the last error of the calling thread is returned by mpx_last_error().
 *   - pointers without suffix are DEVICE pointers on the plan's device; `_h` = host memory.
 *   - all matrices row-major with an explicit leading dimension (in elements).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Device
 *     entry points only enqueue work; the `_host` variants synchronise before returning.
 *   - there is no CPU fallback: without a CUDA device every call fails with MPX_ERR_CUDA.
 */
#ifndef MINTERPY_B200_H
#define MINTERPY_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    MPX_OK = 0,
    MPX_ERR_ARG = 1,        /* bad size / null pointer (reference: ValueError)              */
    MPX_ERR_INCOMPLETE = 2, /* dds on a set that is not complete (core/tree.py:49-54)       */
    MPX_ERR_CUDA = 3,       /* CUDA runtime error, message in mpx_last_error()              */
    MPX_ERR_UNSUPPORTED = 4,/* shape outside the compiled kernel variants   */
    MPX_ERR_NOMEM = 5
};

typedef struct mpx_plan mpx_plan;

/* Library version, major*10000 + minor*100 + patch. */
int mpx_version(void);

/* Message of the calling thread's last failed call ("" if none). */
const char *mpx_last_error(void);

/*
 * Build the immutable, device-resident description of one polynomial space: the
 * exponent matrix of a MultiIndexSet and the generating points of its Grid.
 *   replaces: the per-call preparation in eval_newton_monomials
 *             (src/minterpy/utils.py:218-231: max_exponents, placeholders) and the
 *             MultiIndexTree metadata used by dds (src/minterpy/core/tree.py:47-104,
 *             src/minterpy/dds.py:30-235, 473-602).
 * exponents_h: (N x m) int64 row-major, any row order, rows need not be unique or complete
 *              for evaluation; mpx_dds additionally needs a lexicographically sorted,
 *              complete (downward-closed) set and reports MPX_ERR_INCOMPLETE otherwise.
 * gp_h:        (gp_rows x m) float64 row-major generating points; column i is read at rows
 *              0 .. max_exponent_i (evaluation reads rows < max_exponent_i only, exactly as
 *              src/minterpy/jit_compiled_utils.py:102-112).
 */
int mpx_plan_create(int device, const int64_t *exponents_h, int64_t N, int m,
                    const double *gp_h, int64_t gp_rows, mpx_plan **plan_out);
void mpx_plan_destroy(mpx_plan *plan);

/* Plan properties (host side, no device work). */
enum {
    MPX_INFO_N = 0,            /* number of monomials                                        */
    MPX_INFO_M = 1,            /* spatial dimension                                          */
    MPX_INFO_COMPLETE = 2,     /* 1 if sorted + downward closed (dds allowed)                */
    MPX_INFO_EVAL_FMA = 3,     /* FP64 FMAs the evaluation kernel executes per point and
                                  polynomial (tree form), for roofline accounting            */
    MPX_INFO_DDS_UPDATES = 4,  /* elementary divided-difference updates per value column     */
    MPX_INFO_LEAVES = 5,       /* number of dimension-0 fibers (tree leaves)                 */
    MPX_INFO_DEVICE = 6,
    MPX_INFO_SCRATCH_BYTES = 7,/* device scratch currently held by the plan                  */
    MPX_INFO_EVAL_OTHER_FP64 = 8, /* FP64 adds/multiplies besides the FMAs, per point        */
    MPX_INFO_CHUNKS = 9,       /* chunks of the evaluation program                           */
    MPX_INFO_SMALLK_KERNEL = 10, /* kernel for k below MPX_INFO_MMA_MIN_K: 0 = DFMA tree, 1 = tensor-core tree,
                                  2 = shared-memory-table tensor kernel (long dimension 0),
                                  3 = Horner's rule (m = 1, any k)                           */
    MPX_INFO_TMMA_FMA = 11,    /* FMAs per point and polynomial of the tensor-core tree kernel
                                  (tile padding included), 0 if it was not built             */
    MPX_INFO_UNIFORM_PROGRAM = 13, /* 1 if the tree kernel takes its record stream as a kernel
                                  parameter (uniform control flow; regular index sets)        */
    MPX_INFO_MMA_MIN_K = 12    /* smallest k evaluated by the many-polynomial tensor kernel
                                  (4, 8, 13, 27 or 33, by leaf length; 1: every k; 0: never, m > 8 or m = 1) */
};
int mpx_plan_info(const mpx_plan *plan, int what, int64_t *value_out);

/*
 * Evaluate k polynomials in Newton form at P points.
 *   replaces: eval_newton_polynomials(xx, coefficients, exponents, generating_points)
 *             src/minterpy/utils.py:247-343 (and its batch variant :346-399; the P x N
 *             monomial matrix is never materialised, so batch_size has no meaning here).
 * xx (P x m, ld_xx), coeffs (N x k, ld_c; row j belongs to exponent row j given to
 * mpx_plan_create), out (P x k, ld_out).  P == 0 is a no-op.
 * Uses plan-owned scratch (packed coefficients) that grows on demand: calls on one plan
 * must be issued on one stream at a time.
 */
int mpx_eval(mpx_plan *plan, const double *xx, int64_t P, int64_t ld_xx,
             const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
             void *stream);

/*
 * mpx_eval for callers that must not allocate device memory.  (No counterpart in the
 * reference: the reference has no device memory at all.)
 * Meant for a host program that gave all device memory to its own pool at start-up, where a
 * later cudaMalloc fails, or that cannot take the device-wide synchronisation of
 * cudaMalloc / cudaFree inside its evaluation loop.  Same arguments and results as mpx_eval,
 * but the packed coefficients go to a fixed 8 MiB array inside the library's device image
 * (held by every GPU that loads the library, until exit) instead of plan scratch that may
 * have to grow; the call allocates and frees nothing.
 *   - limit: the packed coefficients of one call must fit 8 MiB.  They take
 *     chunks * 1024 * k * 8 bytes for k below MPX_INFO_MMA_MIN_K (chunks = MPX_INFO_CHUNKS) and
 *     about 1.1 * N * k * 8 bytes (k rounded up to 8, 16, 32 or a multiple of 64) from there on:
 *     up to 32 polynomials per call at BASELINE config 2.  Beyond that the call returns MPX_ERR_UNSUPPORTED and does
 *     nothing -- split k, or use mpx_eval.
 *   - the first call on a device looks the array's address up (set-up); make that first call
 *     before starting any stream capture, and before the application takes all device memory:
 *     with the runtime's default lazy module loading the array is only placed on the GPU at
 *     that moment (or run with CUDA_MODULE_LOADING=EAGER).  A set-up that finds no memory left
 *     returns MPX_ERR_NOMEM.
 *   - the array is shared by all plans and calls on the device and keeps the last call's
 *     contents; each call rewrites every slot it reads, so nothing needs clearing, but two
 *     calls whose kernels overlap on the GPU overwrite each other's coefficients: serialise
 *     them.
 */
int mpx_eval_noalloc(mpx_plan *plan, const double *xx, int64_t P, int64_t ld_xx,
                     const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                     void *stream);

/* Same, host buffers in and out (pinned or pageable): H2D, kernels and D2H are pipelined
 * over blocks of points.  This is what the NumPy path of the Python mirror calls. */
int mpx_eval_host(mpx_plan *plan, const double *xx_h, int64_t P, int64_t ld_xx,
                  const double *coeffs_h, int k, int64_t ld_c, double *out_h,
                  int64_t ld_out);

/*
 * The (P x N) matrix of Newton monomials.
 *   replaces: eval_newton_monomials(x, exponents, generating_points, triangular)
 *             src/minterpy/utils.py:177-244 -> src/minterpy/jit_compiled_utils.py:130-211.
 * triangular != 0 reproduces the reference's lower-triangular mode (:187-199): row p gets
 * monomials 0..p, the rest 0.0.  Results are bit-identical to the reference (same
 * multiplication order, no FMA).
 */
int mpx_monomials(mpx_plan *plan, const double *xx, int64_t P, int64_t ld_xx, double *out,
                  int64_t ld_out, int triangular, void *stream);

/*
 * Multivariate divided differences, in place, on k value columns.
 *   replaces: dds(fct_values, tree)  src/minterpy/dds.py:794-826 (jit_dds :697-791,
 *             project_n_update :642-694, dds_1_dimensional :605-639).
 * vals (N x k, ld): Lagrange coefficients (function values on the unisolvent nodes) in,
 * Newton coefficients out.  The reference works on a copy; the Python mirror makes that
 * copy.  Bit-identical to the reference (true IEEE division, same operation order).
 */
int mpx_dds(mpx_plan *plan, double *vals, int k, int64_t ld, void *stream);

/*
 * Same with flags.  MPX_DDS_MANY_ZEROS: a hint that the value array is mostly exact zeros --
 * columns of the identity, as when the matrix of the Lagrange-to-Newton operator is asked for
 * (array_repr_full, src/minterpy/core/ABC/operator_abstract.py:47-80 over
 * transformations/utils.py:155-179).  Zeros put -0 numerators into every step, which the
 * default instantiation hands to its true-division redo path; with the hint a second
 * instantiation keeps them on the fast path (cfg4, 256 identity columns: 0.62 -> 0.32 ms) at a
 * cost of 8 % on ordinary values.  Results are bit-identical with and without the hint.
 */
#define MPX_DDS_MANY_ZEROS 1
int mpx_dds_ex(mpx_plan *plan, double *vals, int k, int64_t ld, int flags, void *stream);

/* Same, host buffers: out_h (N x k) receives the result, in_h is not modified. */
int mpx_dds_host(mpx_plan *plan, const double *in_h, int k, int64_t ld_in, double *out_h,
                 int64_t ld_out);

/*
 * Dense operator apply out = T @ c.
 *   replaces: MatrixOperator.__matmul__  src/minterpy/schemes/matrix_operator.py:13-22
 * T (rows x cols, ld_t), c (cols x k, ld_c), out (rows x k, ld_out), all device, float64.
 * One or two coefficient columns run as FMA dot products, more as one pass over T per 32
 * columns on the FP64 tensor cores; T is read once per pass and that read bounds the call.
 */
int mpx_apply_dense(int device, const double *T, int64_t rows, int64_t cols, int64_t ld_t,
                    const double *c, int k, int64_t ld_c, double *out, int64_t ld_out,
                    void *stream);

/*
 * The same with flags.  MPX_DENSE_LOWER: T is lower triangular (every Newton<->Lagrange matrix
 * of a sorted index set is: src/minterpy/transformations/utils.py:35-55 builds it with
 * triangular=True, and so is its inverse); entries to the right of the diagonal are taken as
 * zero and mostly not read at all -- about half the bytes and half the time.
 */
enum { MPX_DENSE_LOWER = 1 };
int mpx_apply_dense_ex(int device, const double *T, int64_t rows, int64_t cols, int64_t ld_t,
                       const double *c, int k, int64_t ld_c, double *out, int64_t ld_out,
                       int flags, void *stream);

/*
 * Evaluate a partial derivative of k polynomials in Newton form at P points.
 *   replaces: deriv_newt_eval(x, coefficients, exponents, generating_points,
 *             derivative_order_along)  src/minterpy/polynomials/utils.py:8-112, the evaluation
 *             behind NewtonPolynomial.diff / partial_diff
 *             (src/minterpy/polynomials/newton_polynomial.py:55-86).
 * orders_h: m derivative orders (host, >= 0); all zero is mpx_eval.  Other arguments as
 * mpx_eval; unlike the reference, k > 1 coefficient columns are accepted.  The derivative of a
 * product table entry prod_{t<q}(x - g_t) is formed by the product-rule recurrence (O(q * order)
 * flops) instead of the reference's sum over itertools.combinations, so values agree up to
 * rounding, not bit for bit.  Few polynomials (k below MPX_INFO_MMA_MIN_K) run on the tree kernel
 * with differentiated tables (plan-owned scratch for the tables of dimensions >= 2, one stream
 * at a time like mpx_eval); more on the tensor-core kernel, which reports MPX_ERR_UNSUPPORTED
 * when m > 8 or the per-point tables of the index set do not fit shared memory.
 */
int mpx_eval_deriv(mpx_plan *plan, const double *xx, int64_t P, int64_t ld_xx,
                   const double *coeffs, int k, int64_t ld_c, const int32_t *orders_h,
                   double *out, int64_t ld_out, void *stream);

/*
 * Index-set predicates on a host exponent matrix (N x m, int64, row-major); no device work.
 *   mpx_index_is_sorted:   rows strictly increasing in the reference's lexicographic order
 *                          (last column most significant), i.e. sorted and duplicate-free.
 *   mpx_index_is_complete: replaces is_lexicographically_complete
 *                          src/minterpy/core/utils.py:234-262 (MultiIndexSet.is_complete,
 *                          src/minterpy/core/multi_index.py:120-131): every index that is one
 *                          smaller in one component is in the set.  O(N m^2) instead of the
 *                          reference's O(N^2 m).  Unsorted input is MPX_ERR_ARG.
 */
int mpx_index_is_sorted(const int64_t *exponents_h, int64_t N, int m, int *sorted_out);
int mpx_index_is_complete(const int64_t *exponents_h, int64_t N, int m, int *complete_out);

#ifdef __cplusplus
}
#endif
#endif /* MINTERPY_B200_H */
