/*
 * oracle/newton_oracle.c -- CPU restatement of minterpy's Newton hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under minterpy_b200/ imports, links or calls this
 * file; it exists so that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs have an independent checker for the CUDA kernels.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks every function below against
 * outputs of the unmodified reference (imported from /root/reference/src by
 * tests/golden/make_golden.py, which is committed together with the vectors it wrote to
 * tests/golden/).  eval: <= 1e-13 * max|ref| (the reference ends in a BLAS product whose
 * summation order is unspecified), derivative evaluation likewise; monomials and dds:
 * bit-exact.
 *
 * Each function cites the reference code it restates (paths relative to /root/reference).
 * Arithmetic is plain IEEE double; build with -ffp-contract=off so gcc cannot fuse a*b+c
 * (the reference's Numba/NumPy code never produces an FMA on these expressions).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

/* max_exponents = np.max(exponents, axis=0)            src/minterpy/utils.py:225 */
static void max_exponents(const int64_t *e, int64_t N, int m, int64_t *mx)
{
    for (int i = 0; i < m; ++i) mx[i] = 0;
    for (int64_t j = 0; j < N; ++j)
        for (int i = 0; i < m; ++i)
            if (e[j * m + i] > mx[i]) mx[i] = e[j * m + i];
}

/*
 * One query point: the chained-product table and all N Newton monomials.
 *   src/minterpy/jit_compiled_utils.py:62-126 (eval_newton_monomials_single)
 * products is ((maxmax+1) x m) row-major, row 0 never read (line 96-98 of the reference).
 * n_active < N reproduces the `triangular` branch (lines 187-199): monomials
 * n_active..N-1 are set to 0.0.
 */
static void monomials_single(const double *x, const int64_t *e, int64_t N, int m,
                             const double *gp, int64_t gp_ld, const int64_t *mx,
                             double *products, double *mono, int64_t n_active)
{
    for (int i = 0; i < m; ++i) {              /* :102-112 */
        double prod = 1.0;
        for (int64_t j = 0; j < mx[i]; ++j) {
            prod *= x[i] - gp[j * gp_ld + i];
            products[(j + 1) * m + i] = prod;
        }
    }
    for (int64_t j = 0; j < n_active; ++j) {   /* :115-126 */
        double v = 1.0;
        for (int i = 0; i < m; ++i) {
            int64_t ex = e[j * m + i];
            if (ex > 0) v *= products[ex * m + i];
        }
        mono[j] = v;
    }
    for (int64_t j = n_active; j < N; ++j) mono[j] = 0.0;
}

static int64_t maxmax(const int64_t *mx, int m)
{
    int64_t r = 0;
    for (int i = 0; i < m; ++i) if (mx[i] > r) r = mx[i];
    return r;
}

/*
 * eval_newton_monomials(x, exponents, generating_points, triangular)
 *   src/minterpy/utils.py:177-244 -> jit_compiled_utils.py:130-211
 * xx (P x m), exponents (N x m) int64, gp (gp_rows x m), out (P x N); all row-major.
 */
int orc_eval_newton_monomials(const double *xx, int64_t P, const int64_t *e, int64_t N,
                              int m, const double *gp, int64_t gp_rows, int triangular,
                              double *out)
{
    int64_t mx[64];
    if (m < 1 || m > 64 || N < 1) return 1;
    max_exponents(e, N, m, mx);
    for (int i = 0; i < m; ++i) if (mx[i] > gp_rows) return 2;
    double *products = malloc(sizeof(double) * (size_t)(maxmax(mx, m) + 1) * m);
    if (!products) return 3;
    for (int64_t p = 0; p < P; ++p) {
        int64_t n_active = triangular ? (p + 1 < N ? p + 1 : N) : N;
        monomials_single(xx + p * m, e, N, m, gp, m, mx, products, out + p * N, n_active);
    }
    free(products);
    return 0;
}

/*
 * eval_newton_polynomials(xx, coefficients, exponents, generating_points)
 *   src/minterpy/utils.py:247-343: monomial matrix (P x N), then `@ coefficients` (:332).
 * The P x N matrix is never stored here: each point's monomial row is contracted at once,
 * summing j = 0..N-1 in order (BLAS' own order is unspecified, hence parity by tolerance).
 * coeffs (N x k), out (P x k).  nthreads > 1 splits the points over pthreads (the
 * reference's Numba loop is serial; threads are only for the timed all-cores baseline).
 */
struct poly_job {
    const double *xx, *coeffs, *gp;
    const int64_t *e, *mx;
    double *out;
    int64_t p0, p1, N, tab;
    int k, m, err;
};

static void *poly_worker(void *arg)
{
    struct poly_job *jb = arg;
    const int64_t N = jb->N;
    const int k = jb->k, m = jb->m;
    double *products = malloc(sizeof(double) * (size_t)jb->tab);
    double *mono = malloc(sizeof(double) * (size_t)N);
    if (!products || !mono) {
        jb->err = 3;
    } else {
        for (int64_t p = jb->p0; p < jb->p1; ++p) {
            monomials_single(jb->xx + p * m, jb->e, N, m, jb->gp, m, jb->mx, products,
                             mono, N);
            double *o = jb->out + p * k;
            for (int q = 0; q < k; ++q) o[q] = 0.0;
            for (int64_t j = 0; j < N; ++j) {
                const double mj = mono[j];
                const double *c = jb->coeffs + j * k;
                for (int q = 0; q < k; ++q) o[q] += mj * c[q];
            }
        }
    }
    free(products);
    free(mono);
    return NULL;
}

int orc_eval_newton_polynomials(const double *xx, int64_t P, const double *coeffs, int k,
                                const int64_t *e, int64_t N, int m, const double *gp,
                                int64_t gp_rows, double *out, int nthreads)
{
    int64_t mx[64];
    if (m < 1 || m > 64 || N < 1 || k < 1) return 1;
    max_exponents(e, N, m, mx);
    for (int i = 0; i < m; ++i) if (mx[i] > gp_rows) return 2;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    if ((int64_t)nthreads > P) nthreads = P > 0 ? (int)P : 1;
    struct poly_job jobs[256];
    pthread_t tid[256];
    int err = 0;
    for (int t = 0; t < nthreads; ++t) {
        struct poly_job *jb = &jobs[t];
        jb->xx = xx; jb->coeffs = coeffs; jb->gp = gp; jb->e = e; jb->mx = mx; jb->out = out;
        jb->p0 = P * t / nthreads; jb->p1 = P * (t + 1) / nthreads;
        jb->N = N; jb->tab = (maxmax(mx, m) + 1) * m; jb->k = k; jb->m = m; jb->err = 0;
    }
    for (int t = 1; t < nthreads; ++t)
        if (pthread_create(&tid[t], NULL, poly_worker, &jobs[t]) != 0) {
            poly_worker(&jobs[t]);      /* could not spawn: run it here */
            tid[t] = 0;
        }
    poly_worker(&jobs[0]);
    for (int t = 1; t < nthreads; ++t) if (tid[t]) pthread_join(tid[t], NULL);
    for (int t = 0; t < nthreads; ++t) if (jobs[t].err) err = jobs[t].err;
    return err;
}

/* lexicographic compare, last column most significant: core/utils.py:44-92 (np.lexsort) */
/*
 * deriv_newt_eval(x, coefficients, exponents, generating_points, derivative_order_along)
 *   src/minterpy/polynomials/utils.py:8-112
 * Per point and dimension i with order o = orders[i] > 0 the table entry q becomes
 *   o! * sum over all (q-o)-element subsets S of {0..q-1}, in itertools.combinations order,
 *        of prod_{t in S} (x_i - gp[t, i])                       (:84-93)
 * with entries below o equal to 0 and entry o equal to o! (:73-82; when o is not smaller
 * than the table height nothing else is written, :76-77); dimensions with o == 0 keep the
 * chained products (:61-68).  A monomial whose exponent is 0 in a differentiated dimension
 * is 0 (:104-108).  The result is the dot product with the coefficients (:110; BLAS order is
 * unspecified, summed here j = 0..N-1).  coeffs (N x k) -- the reference takes k = 1 only.
 * The subset sum costs C(q, o) products: keep q and o small in tests.
 */
static double factorial_as_double(int64_t o)
{
    double f = 1.0;                       /* exact up to 22!, beyond that rounded like float(int) */
    for (int64_t t = 2; t <= o; ++t) f *= (double)t;
    return f;
}

/* sum over r-subsets of the first q factors, subsets in lexicographic order */
static double subset_product_sum(const double *factors, int64_t q, int64_t r)
{
    int64_t idx[256];
    double res = 0.0;
    for (int64_t t = 0; t < r; ++t) idx[t] = t;
    for (;;) {
        double prod = 1.0;                /* np.prod of the chosen factors, left to right */
        for (int64_t t = 0; t < r; ++t) prod *= factors[idx[t]];
        res += prod;
        int64_t t = r - 1;
        while (t >= 0 && idx[t] == q - r + t) --t;
        if (t < 0) break;
        ++idx[t];
        for (int64_t u = t + 1; u < r; ++u) idx[u] = idx[u - 1] + 1;
    }
    return res;
}

int orc_deriv_newt_eval(const double *xx, int64_t P, const double *coeffs, int k,
                        const int64_t *e, int64_t N, int m, const double *gp, int64_t gp_rows,
                        const int64_t *orders, double *out)
{
    int64_t mx[64];
    if (m < 1 || m > 64 || N < 1 || k < 1) return 1;
    max_exponents(e, N, m, mx);
    for (int i = 0; i < m; ++i) if (mx[i] > gp_rows || mx[i] > 255 || orders[i] < 0) return 2;
    const int64_t rows = maxmax(mx, m) + 1;            /* num_prods */
    double *products = calloc((size_t)rows * m, sizeof(double));
    double *factors = malloc(sizeof(double) * 256);
    if (!products || !factors) { free(products); free(factors); return 3; }
    for (int64_t p = 0; p < P; ++p) {
        const double *x = xx + p * m;
        for (int i = 0; i < m; ++i) {
            const int64_t o = orders[i];
            if (o == 0) {
                double prod = 1.0;
                for (int64_t j = 0; j < mx[i]; ++j) {
                    prod *= x[i] - gp[j * m + i];
                    products[(j + 1) * m + i] = prod;
                }
                continue;
            }
            for (int64_t j = 0; j < o && j < rows; ++j) products[j * m + i] = 0.0;
            if (o >= rows) continue;
            const double fact = factorial_as_double(o);
            products[o * m + i] = fact;
            for (int64_t q = o + 1; q <= mx[i]; ++q) {
                for (int64_t t = 0; t < q; ++t) factors[t] = x[i] - gp[t * m + i];
                double res = subset_product_sum(factors, q, q - o);
                res *= fact;
                products[q * m + i] = res;
            }
        }
        double *o_row = out + p * k;
        for (int q = 0; q < k; ++q) o_row[q] = 0.0;
        for (int64_t j = 0; j < N; ++j) {
            double v = 1.0;
            for (int i = 0; i < m; ++i) {
                const int64_t ex = e[j * m + i];
                if (ex > 0) v *= products[ex * m + i];
                else if (orders[i] > 0) v = 0.0;
            }
            for (int q = 0; q < k; ++q) o_row[q] += coeffs[j * k + q] * v;
        }
    }
    free(products);
    free(factors);
    return 0;
}

static int lex_cmp(const int64_t *a, const int64_t *b, int m)
{
    for (int i = m - 1; i >= 0; --i) {
        if (a[i] < b[i]) return -1;
        if (a[i] > b[i]) return 1;
    }
    return 0;
}

static int64_t find_row(const int64_t *e, int64_t N, int m, const int64_t *key)
{
    int64_t lo = 0, hi = N - 1;
    while (lo <= hi) {
        int64_t mid = (lo + hi) / 2;
        int c = lex_cmp(e + mid * m, key, m);
        if (c == 0) return mid;
        if (c < 0) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

/*
 * Successor map along dimension d: succ[j] = row of exponents[j] + unit_d, or -1.
 * Also checks what core/tree.py:49-54 demands (a complete = downward-closed,
 * lexicographically sorted set): every row with e_d > 0 must have its predecessor.
 * Returns 0, or 4 if the set is not sorted / not complete.
 */
static int build_succ(const int64_t *e, int64_t N, int m, int d, int64_t *succ)
{
    int64_t key[64];
    for (int64_t j = 0; j < N; ++j) {
        memcpy(key, e + j * m, sizeof(int64_t) * m);
        key[d] += 1;
        succ[j] = find_row(e, N, m, key);
        if (e[j * m + d] > 0) {
            key[d] -= 2;
            if (find_row(e, N, m, key) < 0) return 4;
        }
    }
    return 0;
}

/*
 * dds(fct_values, tree)      src/minterpy/dds.py:794-826 -> jit_dds :697-791
 *
 * Restated in fiber form.  jit_dds walks dim_par = m-1..1; for each parent node and each
 * ordered pair of children l < r it does (project_n_update, :642-694)
 *     v_r[:] = (v_r - v_l[mask]) / (gp[e_l + (r-l), dim] - gp[e_l, dim])
 * and finally a 1-D scheme on every leaf (dds_1_dimensional, :605-639).  `mask` pairs
 * each entry of v_r with the entry of v_l that has the same lower-dimensional exponents,
 * i.e. the two lie on one axis-parallel fiber along `dim`; children of a node carry
 * exponents 0,1,2,.. so e_l = l.  Per element this is exactly, for d = m-1..0 and every
 * fiber F along d (rows that differ only in e_d = 0..len-1):
 *     for i = 1..len-1: for r = i..len-1: v[r] = (v[r] - v[i-1]) / (g[r] - g[i-1])
 * with g = gp[:, d] -- same operands, same order of the two subtractions and the division
 * for every element, hence bit-identical (pinned by the golden vectors).
 * vals (N x k) row-major is overwritten (the reference works on a copy, :809).
 */
int orc_dds(double *vals, int64_t N, int k, const int64_t *e, int m, const double *gp,
            int64_t gp_rows)
{
    if (m < 1 || m > 64 || N < 1 || k < 1) return 1;
    int64_t mx[64];
    max_exponents(e, N, m, mx);
    for (int i = 0; i < m; ++i) if (mx[i] >= gp_rows) return 2;
    for (int64_t j = 1; j < N; ++j)
        if (lex_cmp(e + (j - 1) * m, e + j * m, m) >= 0) return 4;
    int64_t *succ = malloc(sizeof(int64_t) * (size_t)N);
    int64_t *fiber = malloc(sizeof(int64_t) * (size_t)(maxmax(mx, m) + 1));
    if (!succ || !fiber) { free(succ); free(fiber); return 3; }
    int rc = 0;
    for (int d = m - 1; d >= 0 && rc == 0; --d) {
        rc = build_succ(e, N, m, d, succ);
        if (rc) break;
        for (int64_t j0 = 0; j0 < N; ++j0) {
            if (e[j0 * m + d] != 0) continue;
            int64_t len = 0;
            for (int64_t j = j0; j >= 0; j = succ[j]) fiber[len++] = j;
            for (int64_t i = 1; i < len; ++i) {
                const double gl = gp[(i - 1) * m + d];
                const double *vl = vals + fiber[i - 1] * k;
                for (int64_t r = i; r < len; ++r) {
                    const double diff = gp[r * m + d] - gl;
                    double *vr = vals + fiber[r] * k;
                    for (int q = 0; q < k; ++q) vr[q] = (vr[q] - vl[q]) / diff;
                }
            }
        }
    }
    free(succ);
    free(fiber);
    return rc;
}

int orc_num_threads(void)
{
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}
