// K3: multivariate divided differences (Lagrange -> Newton), fiber form.
//   replaces dds / jit_dds / project_n_update / dds_1_dimensional,
//   src/minterpy/dds.py:794-826, 697-791, 642-694, 605-639.
//
// For d = m-1 .. 0, every fiber along dimension d (rows that differ only in e_d = 0..len-1)
// gets the 1-D scheme  for i=1..len-1: for r=i..len-1: v[r] = (v[r]-v[i-1]) / (g[r]-g[i-1]).
// This is, element by element, the same sequence of IEEE operations as the reference's
// tree walk (see oracle/newton_oracle.c:orc_dds for the argument), so results are
// bit-identical: true division, no FMA contraction possible in (a-b)/c.
// One thread owns one (fiber, column) pair and keeps the fiber in registers; the k value
// columns are the fastest thread index, so global accesses are coalesced along a row.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "mpx_common.h"
#include "mpx_ptx.h"

namespace {

template <int LMAX>
__global__ void __launch_bounds__(256)
dds_pass_kernel(double *__restrict__ vals, int64_t ld, int k, const int32_t *__restrict__ off,
                const int32_t *__restrict__ rows, int n_fibers, const double *__restrict__ diff,
                int L)
{
    __shared__ double sdiff[LMAX * LMAX];
    for (int i = threadIdx.x; i < L * L; i += blockDim.x) {
        int l = i / L, r = i % L;
        sdiff[l * LMAX + r] = diff[i];
    }
    __syncthreads();
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int col = (int)(t % k);
    const int64_t f = t / k;
    if (f >= n_fibers) return;
    const int o = off[f];
    const int len = off[f + 1] - o;
    if (len < 2) return;
    double v[LMAX];
    double *ptr[LMAX];
#pragma unroll
    for (int r = 0; r < LMAX; ++r)
        if (r < len) {
            ptr[r] = vals + (int64_t)rows[o + r] * ld + col;
            v[r] = *ptr[r];
        }
#pragma unroll
    for (int i = 1; i < LMAX; ++i) {
        if (i < len) {
#pragma unroll
            for (int r = i; r < LMAX; ++r)
                if (r < len) v[r] = __ddiv_rn(__dsub_rn(v[r], v[i - 1]), sdiff[(i - 1) * LMAX + r]);
        }
    }
#pragma unroll
    for (int r = 1; r < LMAX; ++r)
        if (r < len) *ptr[r] = v[r];
}

// any fiber length: the fiber sits in global memory (L1/L2), row-oriented sweep
__global__ void __launch_bounds__(256)
dds_pass_generic_kernel(double *__restrict__ vals, int64_t ld, int k,
                        const int32_t *__restrict__ off, const int32_t *__restrict__ rows,
                        int n_fibers, const double *__restrict__ diff, int L)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int col = (int)(t % k);
    const int64_t f = t / k;
    if (f >= n_fibers) return;
    const int o = off[f];
    const int len = off[f + 1] - o;
    for (int r = 1; r < len; ++r) {
        double *pr = vals + (int64_t)rows[o + r] * ld + col;
        double x = *pr;
        for (int i = 1; i <= r; ++i) {
            const double vl = vals[(int64_t)rows[o + i - 1] * ld + col];
            x = __ddiv_rn(__dsub_rn(x, vl), diff[(i - 1) * L + r]);
        }
        *pr = x;
    }
}


// ---- fused kernel: one CTA = one slice x 32 value columns -------------------------------
// The slice (rows that agree outside the dimension group) is gathered into shared memory as
// v[row][column]; then, for d = hi .. lo, every fiber of dimension d inside the slice gets the
// 1-D scheme, one warp per fiber, lane = column:
//   for i = 1..len-1: for r = i..len-1: v[r] = (v[r] - v[i-1]) / (g[r] - g[i-1]).
//
// Division.  Every divisor comes from a small table, so its correctly rounded reciprocal
// y = RN(1/d) is tabulated too (host division) and the quotient is formed as
//   q = a*y;  r = fma(-d, q, a);  q' = fma(r, y, q)
// -- Markstein's correction step, which returns the IEEE quotient RN(a/d) when y is the
// correctly rounded reciprocal, d's significand is not all ones and nothing over/underflows.
// The plan tabulates reciprocal 0 for divisors that do not qualify; the kernel falls back to
// __ddiv_rn for those and for numerators outside [2^-900, 2^900] (zero, subnormal, huge, inf,
// nan).  tools/div_check.cu
// compares the sequence with __ddiv_rn on 1e10 random and near-midpoint cases: no mismatch.
struct DdsDimArgs {
    const int32_t *slice_fib0;
    const int32_t *fib_off;
    const uint16_t *fib_rows;
    const double *diff;
    const double *recip;
    const double *gcol;
    int L;
};
struct DdsGroupArgs {
    const int32_t *slice_off;
    const int32_t *slice_rows;
    int n_dims;
    int max_rows;
    int max_fibers[MPX_MAX_M];
    DdsDimArgs dim[MPX_MAX_M];
};

constexpr int DDS_WARPS = 8;
constexpr int DDS_MAX_FIBER = 256;

// out of line on purpose: inlined, the compiler evaluates the whole division speculatively
// next to the fast sequence
__device__ __noinline__ double dds_true_division(double a, double d) { return __ddiv_rn(a, d); }

template <bool FAST>
__device__ __forceinline__ double dds_update(double x, double pv, double d, double y)
{
    const double a = __dsub_rn(x, pv);
    if constexpr (FAST) {
        const double q = __dmul_rn(a, y);
        const double r = __fma_rn(-d, q, a);
        const double q2 = __fma_rn(r, y, q);
        const unsigned ex = ((unsigned)__double2hiint(a) >> 20) & 0x7ffu;
        // 2^-900 <= |a| < 2^900 and the divisor qualified (its tabulated reciprocal is not 0)
        if (ex - 123u < 1800u && __double2hiint(y) != 0) return q2;
        // +-0 over a qualified divisor: the zero a * y is the quotient (sign of a x sign of d)
        if ((((unsigned)__double2hiint(a) & 0x7fffffffu) | (unsigned)__double2loint(a)) == 0u && __double2hiint(y) != 0)
            return q;
        return dds_true_division(a, d);
    } else {
        return __ddiv_rn(a, d);       // here d = g[r] - g[i-1] was formed by the caller
    }
}

template <bool FAST>
__global__ void __launch_bounds__(DDS_WARPS * 32)
dds_group_kernel(double *__restrict__ vals, int64_t ld, int k, const DdsGroupArgs g)
{
    // shared memory: v[max_rows][32] | per dim: diff (L*L), recip (L*L)  [FAST]  or g (L) |
    // per dim: fiber
    // offsets (max_fibers + 1) and fiber rows as byte offsets into v (max_rows), both u32
    extern __shared__ __align__(16) unsigned char dsm[];
    double *v = reinterpret_cast<double *>(dsm);
    double *stab = v + (size_t)g.max_rows * MPX_DDS_COLS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int s = blockIdx.x;
    const int col = blockIdx.y * MPX_DDS_COLS + lane;
    const bool active = col < k;
    const int r0 = g.slice_off[s];
    const int nrows = g.slice_off[s + 1] - r0;
    size_t toff = 0;
    for (int t = 0; t < g.n_dims; ++t) {
        if constexpr (FAST) {
            const int LL = g.dim[t].L * g.dim[t].L;
            for (int i = threadIdx.x; i < LL; i += blockDim.x) {
                stab[toff + i] = g.dim[t].diff[i];
                stab[toff + LL + i] = g.dim[t].recip[i];
            }
            toff += 2 * LL;
        } else {
            // long fibers: no room for L x L tables -- keep the generating points and form
            // every divisor g[r] - g[i-1] on the fly (the very subtraction the host tabulates)
            for (int i = threadIdx.x; i < g.dim[t].L; i += blockDim.x) stab[toff + i] = g.dim[t].gcol[i];
            toff += g.dim[t].L;
        }
    }
    uint32_t *meta = reinterpret_cast<uint32_t *>(stab + toff);
    {
        // fiber tables of this slice: the sweeps below then touch shared memory only
        uint32_t *mp = meta;
        for (int t = 0; t < g.n_dims; ++t) {
            const DdsDimArgs &gd = g.dim[t];
            const int f0 = gd.slice_fib0[s], nf = gd.slice_fib0[s + 1] - f0;
            const int o0 = gd.fib_off[f0];
            for (int i = threadIdx.x; i <= nf; i += blockDim.x) mp[i] = (uint32_t)(gd.fib_off[f0 + i] - o0);
            uint32_t *rows = mp + g.max_fibers[t] + 1;
            for (int i = threadIdx.x; i < nrows; i += blockDim.x)
                rows[i] = (uint32_t)gd.fib_rows[o0 + i] * (uint32_t)(MPX_DDS_COLS * sizeof(double));
            mp = rows + g.max_rows;
        }
    }
    if (active) {
        int r = warp;
        for (; r + 3 * DDS_WARPS < nrows; r += 4 * DDS_WARPS) {       // four gathers in flight
            const int i0 = g.slice_rows[r0 + r], i1 = g.slice_rows[r0 + r + DDS_WARPS];
            const int i2 = g.slice_rows[r0 + r + 2 * DDS_WARPS], i3 = g.slice_rows[r0 + r + 3 * DDS_WARPS];
            const double x0 = vals[(int64_t)i0 * ld + col], x1 = vals[(int64_t)i1 * ld + col];
            const double x2 = vals[(int64_t)i2 * ld + col], x3 = vals[(int64_t)i3 * ld + col];
            v[(size_t)r * MPX_DDS_COLS + lane] = x0;
            v[(size_t)(r + DDS_WARPS) * MPX_DDS_COLS + lane] = x1;
            v[(size_t)(r + 2 * DDS_WARPS) * MPX_DDS_COLS + lane] = x2;
            v[(size_t)(r + 3 * DDS_WARPS) * MPX_DDS_COLS + lane] = x3;
        }
        for (; r < nrows; r += DDS_WARPS)
            v[(size_t)r * MPX_DDS_COLS + lane] = vals[(int64_t)g.slice_rows[r0 + r] * ld + col];
    }
    __syncthreads();
    unsigned char *vlane = reinterpret_cast<unsigned char *>(v + lane);
    toff = 0;
    const uint32_t *mp = meta;
    for (int t = 0; t < g.n_dims; ++t) {
        const DdsDimArgs &gd = g.dim[t];
        const int L = gd.L;
        const double *df = stab + toff;          // FAST: diff table, else the g column
        const double *ry = df + L * L;           // FAST only
        const int nf = gd.slice_fib0[s + 1] - gd.slice_fib0[s];
        const uint32_t *foff = mp;
        const uint32_t *frows = mp + g.max_fibers[t] + 1;
        if (active) {
            for (int f = warp; f < nf; f += DDS_WARPS) {
                const uint32_t *fidx = frows + foff[f];
                const int len = (int)(foff[f + 1] - foff[f]);
                // column-wise sweep: for a fixed i the updates of r = i..len-1 are independent,
                // four of them are kept in flight
                for (int i = 1; i < len; ++i) {
                    const double pv = *reinterpret_cast<double *>(vlane + fidx[i - 1]);
                    if constexpr (FAST) {
                        const double *dfi = df + (i - 1) * L, *ryi = ry + (i - 1) * L;
                        int r = i;
                        for (; r + 3 < len; r += 4) {
                            double *p0 = reinterpret_cast<double *>(vlane + fidx[r]);
                            double *p1 = reinterpret_cast<double *>(vlane + fidx[r + 1]);
                            double *p2 = reinterpret_cast<double *>(vlane + fidx[r + 2]);
                            double *p3 = reinterpret_cast<double *>(vlane + fidx[r + 3]);
                            const double x0 = *p0, x1 = *p1, x2 = *p2, x3 = *p3;
                            const double y0 = dds_update<true>(x0, pv, dfi[r], ryi[r]);
                            const double y1 = dds_update<true>(x1, pv, dfi[r + 1], ryi[r + 1]);
                            const double y2 = dds_update<true>(x2, pv, dfi[r + 2], ryi[r + 2]);
                            const double y3 = dds_update<true>(x3, pv, dfi[r + 3], ryi[r + 3]);
                            *p0 = y0; *p1 = y1; *p2 = y2; *p3 = y3;
                        }
                        for (; r < len; ++r) {
                            double *p0 = reinterpret_cast<double *>(vlane + fidx[r]);
                            *p0 = dds_update<true>(*p0, pv, dfi[r], ryi[r]);
                        }
                    } else {
                        const double gl = df[i - 1];
                        int r = i;
                        for (; r + 1 < len; r += 2) {
                            double *p0 = reinterpret_cast<double *>(vlane + fidx[r]);
                            double *p1 = reinterpret_cast<double *>(vlane + fidx[r + 1]);
                            const double x0 = *p0, x1 = *p1;
                            const double y0 = dds_update<false>(x0, pv, __dsub_rn(df[r], gl), 0.0);
                            const double y1 = dds_update<false>(x1, pv, __dsub_rn(df[r + 1], gl), 0.0);
                            *p0 = y0; *p1 = y1;
                        }
                        for (; r < len; ++r) {
                            double *p0 = reinterpret_cast<double *>(vlane + fidx[r]);
                            *p0 = dds_update<false>(*p0, pv, __dsub_rn(df[r], gl), 0.0);
                        }
                    }
                }
            }
        }
        toff += FAST ? 2 * L * L : L;
        mp = frows + g.max_rows;
        __syncthreads();
    }
    if (active)
        for (int r = warp; r < nrows; r += DDS_WARPS)
            vals[(int64_t)g.slice_rows[r0 + r] * ld + col] = v[(size_t)r * MPX_DDS_COLS + lane];
}

// ---- one warp per fiber, one pass per dimension ------------------------------------------
// A warp gathers ONE fiber (len rows x 32 value columns) into its private piece of shared
// memory, runs the 1-D scheme there with direct addressing (row r of the fiber is row r of the
// buffer), and writes rows 1..len-1 back.  No lane ever touches another lane's column, so the
// kernel has no barrier after the table is staged: warps come and go independently and the
// gather latency of one hides behind the arithmetic of the others -- where the fused kernel
// above spends 40 % of its warp time at barriers and on its gather.  The price is one read and
// one write of the value array per dimension instead of one per dimension group.

template <bool FAST, int FW>
__global__ void __launch_bounds__(FW * 32)
dds_fiber_kernel(double *__restrict__ vals, int64_t ld, int k, const int32_t *__restrict__ woff,
                 const int32_t *__restrict__ wrows, int n_work, const double *__restrict__ dy,
                 const double *__restrict__ gcol, int L)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    // FAST: (divisor, reciprocal) pairs, L*L of them; else the L generating points
    double *tab = reinterpret_cast<double *>(fsm);
    const int tab_doubles = FAST ? 2 * L * L : ((L + 1) & ~1);
    if constexpr (FAST) {
        for (int i = threadIdx.x; i < 2 * L * L; i += blockDim.x) tab[i] = dy[i];
    } else {
        for (int i = threadIdx.x; i < L; i += blockDim.x) tab[i] = gcol[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = blockIdx.x * FW + warp;
    const int col = blockIdx.y * MPX_DDS_COLS + lane;
    if (f >= n_work || col >= k) return;
    const int o = woff[f];
    const int len = woff[f + 1] - o;
    double *b = tab + tab_doubles + (size_t)warp * L * MPX_DDS_COLS + lane;     // b[r * 32]
    const int32_t *rows = wrows + o;
    // (fetching the row numbers once per warp and handing them round by shuffles was measured:
    // no gain, the index loads hit L1)
    {
        int r = 0;
        for (; r + 3 < len; r += 4) {                 // four row segments in flight
            double x[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) x[q] = vals[(int64_t)rows[r + q] * ld + col];
#pragma unroll
            for (int q = 0; q < 4; ++q) b[(r + q) * MPX_DDS_COLS] = x[q];
        }
        for (; r < len; ++r) b[r * MPX_DDS_COLS] = vals[(int64_t)rows[r] * ld + col];
    }
    if constexpr (FAST) {
        // Row-blocked sweep.  Element r receives its updates i = 1..r in order, each using the
        // FINAL value of element i-1, so the rows can be finished one after the other (the same
        // operations on every element as the column-wise order of the reference).  Four rows
        // advance together: they share the read of v[i-1], their four (divisor, reciprocal)
        // pairs are neighbours in the table, and the running values stay in registers -- no
        // store per update, no remainder loop per step.  The last three updates of rows
        // r+1..r+3 use the rows of the block itself.
        const double2 *dy2 = reinterpret_cast<const double2 *>(tab);
        // the (len - 1) % 4 rows that do not fill a block are the FIRST ones: their chains are
        // the shortest
        int r = 1;
        const int first_block = 1 + ((len - 1) & 3);
        for (; r < first_block; ++r) {
            double x = b[r * MPX_DDS_COLS];
            const double2 *t = dy2 + r;
            const double *vp = b;
#pragma unroll 1
            for (int i = 1; i <= r; ++i) {
                const double2 t0 = t[0];
                x = dds_update<true>(x, *vp, t0.x, t0.y);
                t += L;
                vp += MPX_DDS_COLS;
            }
            b[r * MPX_DDS_COLS] = x;
        }
        for (; r + 3 < len; r += 4) {
            double *p = b + r * MPX_DDS_COLS;
            double x0 = p[0], x1 = p[MPX_DDS_COLS], x2 = p[2 * MPX_DDS_COLS], x3 = p[3 * MPX_DDS_COLS];
            const double2 *t = dy2 + r;               // entries (i-1, r..r+3), i-1 = 0
            const double *vp = b;
#pragma unroll 1
            for (int i = 1; i <= r; ++i) {
                const double pv = *vp;
                const double2 t0 = t[0], t1 = t[1], t2 = t[2], t3 = t[3];
                x0 = dds_update<true>(x0, pv, t0.x, t0.y);
                x1 = dds_update<true>(x1, pv, t1.x, t1.y);
                x2 = dds_update<true>(x2, pv, t2.x, t2.y);
                x3 = dds_update<true>(x3, pv, t3.x, t3.y);
                t += L;
                vp += MPX_DDS_COLS;
            }
            {
                const double2 t1 = t[1], t2 = t[2], t3 = t[3];          // table row r
                x1 = dds_update<true>(x1, x0, t1.x, t1.y);
                x2 = dds_update<true>(x2, x0, t2.x, t2.y);
                x3 = dds_update<true>(x3, x0, t3.x, t3.y);
                t += L;
                const double2 u2 = t[2], u3 = t[3];                      // table row r + 1
                x2 = dds_update<true>(x2, x1, u2.x, u2.y);
                x3 = dds_update<true>(x3, x1, u3.x, u3.y);
                t += L;
                const double2 w3 = t[3];                                 // table row r + 2
                x3 = dds_update<true>(x3, x2, w3.x, w3.y);
            }
            p[0] = x0;
            p[MPX_DDS_COLS] = x1;
            p[2 * MPX_DDS_COLS] = x2;
            p[3 * MPX_DDS_COLS] = x3;
        }
    } else {
        for (int i = 1; i < len; ++i) {
            const double pv = b[(i - 1) * MPX_DDS_COLS];
            const double gl = tab[i - 1];
            int r = i;
            for (; r + 1 < len; r += 2) {
                double *p = b + r * MPX_DDS_COLS;
                const double x0 = p[0], x1 = p[MPX_DDS_COLS];
                const double y0 = dds_update<false>(x0, pv, __dsub_rn(tab[r], gl), 0.0);
                const double y1 = dds_update<false>(x1, pv, __dsub_rn(tab[r + 1], gl), 0.0);
                p[0] = y0;
                p[MPX_DDS_COLS] = y1;
            }
            for (; r < len; ++r)
                b[r * MPX_DDS_COLS] = dds_update<false>(b[r * MPX_DDS_COLS], pv, __dsub_rn(tab[r], gl), 0.0);
        }
    }
    for (int r = 1; r < len; ++r)                      // element 0 of a fiber never changes
        vals[(int64_t)rows[r] * ld + col] = b[r * MPX_DDS_COLS];
}


// ---- one warp per fiber, several value columns per lane ------------------------------------
// Same scheme as dds_fiber_kernel, but a lane owns C = 2H value columns (H pairs of neighbouring
// columns, 16-byte accesses everywhere).  On this GPU a shared-memory load costs the FP64 pipe
// about as much as two to three FP64 instructions (tools/lds_dfma_probe.cu), and an update is
// only four of them: with one column per lane the (divisor, reciprocal) pair and the pivot value
// cost more than the arithmetic.  Here one pair serves C updates and one pivot load two, index
// loads and address arithmetic are shared by C columns, and the rare cases the fast division
// cannot take no longer branch inside the sweep: every update computes the fast quotient and ORs
// "not valid" into one sticky flag per lane; a flagged lane redoes its columns with true
// divisions from the values still unchanged in global memory (dds_redo_columns).
// Buffer layout: row r, pair h of the lane at double2 index (r * H + h) * 32 + lane.
// One update of the redo path: the fast division where it is exact (numerator in range, or a
// zero of either sign), the true division by the true divisor otherwise.
__device__ __forceinline__ double dds_redo_one(double x, double pv, double2 t, bool inc,
                                               const double *__restrict__ gcol, int r, int i)
{
    const double a = __dsub_rn(x, pv);
    const double q = __dmul_rn(a, t.y);
    const double rem = __fma_rn(-t.x, q, a);
    double q2 = __fma_rn(rem, t.y, q);
    const int hi = __double2hiint(a);
    if (((unsigned)(hi & 0x7ff00000) - (123u << 20)) < (1800u << 20)) {
        if (inc) q2 = __longlong_as_double(__double_as_longlong(q2) + 1);
        return q2;
    }
    if (((hi & 0x7fffffff) | __double2loint(a)) == 0) return q;    // +-0 / d: the zero a * y
    return dds_true_division(a, __dsub_rn(gcol[r], gcol[i - 1]));  // the subtraction the host tabulated
}

template <int H>
__device__ __noinline__ void dds_redo_columns(double2 *__restrict__ bl, const double *__restrict__ vals,
                                              int64_t ld, int64_t col0, unsigned active,
                                              const int32_t *__restrict__ rows, int len,
                                              const double *__restrict__ gcol,
                                              const double2 *__restrict__ dy2,
                                              const unsigned long long *__restrict__ fm, int L)
{
    for (int r = 0; r < len; ++r)
#pragma unroll
        for (int h = 0; h < H; ++h)
            if (active >> h & 1)
                bl[(r * H + h) * 32] = *reinterpret_cast<const double2 *>(vals + (int64_t)rows[r] * ld + col0 + 64 * h);
    for (int i = 1; i < len; ++i)
        for (int r = i; r < len; ++r) {
            const double2 t = dy2[(i - 1) * L + r];
            const bool inc = (fm[i - 1] >> r) & 1ull;
#pragma unroll
            for (int h = 0; h < H; ++h) {
                double2 x = bl[(r * H + h) * 32];
                const double2 pv = bl[((i - 1) * H + h) * 32];
                x.x = dds_redo_one(x.x, pv.x, t, inc, gcol, r, i);
                x.y = dds_redo_one(x.y, pv.y, t, inc, gcol, r, i);
                bl[(r * H + h) * 32] = x;
            }
        }
}

// Division by a table divisor d with y = RN(1/d):  q = a y;  r = fma(-d, q, a);  q' = fma(r, y, q)
// is the IEEE quotient unless d's significand is all ones (see the fused kernel above).  Such a
// divisor is d = s 2^j (1 - 2^-53), s = +-1 (between Chebyshev points: 1 - 6e-17), and then
//     a / d = A (1 + 2^-53 + 2^-106 + ...),   A = a / (s 2^j)   (exact),
// lies strictly between |A| + ulp/2 and |A| + ulp for every normal A: the quotient is the next
// double after A, away from zero.  The plan stores (s 2^j, s 2^-j) for these entries, with which
// the same three instructions return exactly A, and flags them (INC); one integer increment of
// the bit pattern finishes the division.  tools/div_check.cu compares both forms with __ddiv_rn.
// `bad` collects what neither form may take: a outside [2^-900, 2^900) other than +0.
template <bool FLAGGED, bool ZSAFE = false>
__device__ __forceinline__ double dds_update_flag(double x, double pv, double d, double y, bool inc,
                                                  bool &bad)
{
    const double a = __dsub_rn(x, pv);
    const double q = __dmul_rn(a, y);
    const double r = __fma_rn(-d, q, a);
    double q2 = __fma_rn(r, y, q);
    const int hi = __double2hiint(a);
    const bool in_range = ((unsigned)(hi & 0x7ff00000) - (123u << 20)) < (1800u << 20);
    if constexpr (ZSAFE) {
        // the instantiation for value arrays known to be full of zeros (MPX_DDS_MANY_ZEROS):
        // +-0 / d is the zero a * y, which has the right sign
        const bool zero = ((hi & 0x7fffffff) | __double2loint(a)) == 0;
        bad = bad || !(in_range || zero);
        if constexpr (FLAGGED) {
            if (inc && !zero)
                q2 = __longlong_as_double(__double_as_longlong(q2) + 1);
        }
        return zero ? q : q2;
    }
    const bool plus_zero = (hi | __double2loint(a)) == 0;
    bad = bad || !(in_range || plus_zero);
    if constexpr (FLAGGED) {
        if (inc && !plus_zero)
            q2 = __longlong_as_double(__double_as_longlong(q2) + 1);
    }
    // (-0 is left to the redo path: the correction steps return +0 for -0 / d, d > 0.  Taking
    // the zero a * y instead -- it has the right sign -- was measured: the select costs the
    // ordinary path 8 % (0.301 -> 0.325 ms at cfg4), a select on the high word alone 20 %.)
    return q2;
}

template <int FW, int H, bool ZSAFE>
__global__ void __launch_bounds__(FW * 32)
dds_fiber_cols_kernel(double *__restrict__ vals, int64_t ld, int k, const int32_t *__restrict__ woff,
                      const int32_t *__restrict__ wrows, int n_work, const double *__restrict__ dy,
                      const unsigned long long *__restrict__ fmask, const double *__restrict__ gcol, int L)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    double2 *dy2 = reinterpret_cast<double2 *>(fsm);                  // (divisor, reciprocal), L x L
    // flag words behind the fiber buffers and the barriers: bit r of word l = entry (l, r) flagged
    unsigned long long *fm = reinterpret_cast<unsigned long long *>(dy2 + L * L + (size_t)FW * L * H * 32) + FW;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = blockIdx.x * FW + warp;
    const bool valid = f < n_work;             // (a warp without a fiber still stages the table)
    constexpr int C = 2 * H;
    const int64_t col0 = (int64_t)blockIdx.y * (32 * C) + 2 * lane;      // pair h starts at col0 + 64 h
    // pairs of this lane that exist (k may end inside the warp's column block); a lane without
    // any still takes part: it hands rows to the TMA and sweeps zeros
    unsigned active = 0;
#pragma unroll
    for (int h = 0; h < H; ++h) active |= (col0 + 64 * h < k) ? 1u << h : 0u;
    // `wrows` holds L row numbers per fiber (padded): the lane's rows for the gather (row `lane`)
    // and for the write-back (row `lane + 1`) are read while the fiber's length is on its way
    const int32_t *rows = wrows + (size_t)(valid ? f : 0) * L;
    const int32_t row_in = rows[lane < L ? lane : L - 1];
    const int32_t row_out = rows[lane + 1 < L ? lane + 1 : L - 1];
    const int len = valid ? woff[f + 1] - woff[f] : 0;
    double2 *bw = dy2 + L * L + (size_t)warp * L * H * 32;               // this warp's fiber buffer
    double2 *bl = bw + lane;                                             // this lane's slots
    // Gather: row r of the fiber is ONE contiguous piece of global memory (the warp's 64 H value
    // columns), and the buffer keeps it contiguous too, so lane r hands the whole row to the TMA
    // (cp.async.bulk, completion counted on the warp's mbarrier): every row of the fiber is in
    // flight at once, no registers and no address arithmetic per element.
    const int64_t cw0 = (int64_t)blockIdx.y * (32 * C);                  // first column of the warp
    const int ncols = (int)((k - cw0 < 32 * C) ? (k - cw0) : 32 * C);    // even: k is
    const uint32_t row_bytes = (uint32_t)ncols * 8u;
    uint64_t *bar = reinterpret_cast<uint64_t *>(dy2 + L * L + (size_t)FW * L * H * 32) + warp;
    if (lane == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (ncols < 32 * C)                                  // ragged last column block: the sweep
        for (int r = 0; r < len; ++r)                    // must not chew on stale shared memory
#pragma unroll
            for (int h = 0; h < H; ++h) bl[(r * H + h) * 32] = make_double2(0.0, 0.0);
    if (ncols < 32 * C) fence_proxy_async_smem();
    __syncwarp();
    if (valid) {
        if (lane == 0) mbar_arrive_expect_tx(bar, row_bytes * (uint32_t)len);
        if (lane < len)
            bulk_g2s(bw + (size_t)lane * H * 32, vals + (int64_t)row_in * ld + cw0, row_bytes, bar);
        for (int r = lane + 32; r < len; r += 32)
            bulk_g2s(bw + (size_t)r * H * 32, vals + (int64_t)rows[r] * ld + cw0, row_bytes, bar);
    }
    // the tables are staged while the rows are on their way
    for (int i = threadIdx.x; i < L * L; i += blockDim.x) dy2[i] = reinterpret_cast<const double2 *>(dy)[i];
    for (int i = threadIdx.x; i < L; i += blockDim.x) fm[i] = fmask[i];
    __syncthreads();
    if (!valid) return;
    mbar_wait(bar, 0);
    bool bad = false;
    // Row-blocked sweep (see dds_fiber_kernel): element r receives its updates i = 1..r in order,
    // each with the final value of element i-1; four rows advance together.
#ifdef MPX_DDS_PROBE_NO_SWEEP     // probe build: gather and scatter only
    int r = len;
#else
    int r = 1;
#endif
    const int first_block = 1 + ((len - 1) & 3);
    for (; r < first_block; ++r) {
        double2 x[H];
#pragma unroll
        for (int h = 0; h < H; ++h) x[h] = bl[(r * H + h) * 32];
        const double2 *t = dy2 + r;
        const double2 *vp = bl;
#pragma unroll 1
        for (int i = 1; i <= r; ++i) {
            const double2 t0 = t[0];
            const bool inc = (fm[i - 1] >> r) & 1ull;
#pragma unroll
            for (int h = 0; h < H; ++h) {
                const double2 pv = vp[h * 32];
                x[h].x = dds_update_flag<true, ZSAFE>(x[h].x, pv.x, t0.x, t0.y, inc, bad);
                x[h].y = dds_update_flag<true, ZSAFE>(x[h].y, pv.y, t0.x, t0.y, inc, bad);
            }
            t += L;
            vp += H * 32;
        }
#pragma unroll
        for (int h = 0; h < H; ++h) bl[(r * H + h) * 32] = x[h];
    }
    for (; r + 3 < len; r += 4) {
        double2 x[4][H];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int h = 0; h < H; ++h) x[q][h] = bl[((r + q) * H + h) * 32];
        const double2 *t = dy2 + r;               // entries (i-1, r..r+3), i-1 = 0
        const double2 *vp = bl;
#pragma unroll 1
        for (int i = 1; i <= r; ++i) {
            double2 tq[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) tq[q] = t[q];
            const unsigned flags = (unsigned)(fm[i - 1] >> r) & 15u;     // uniform: table data
            if (flags == 0) {
#pragma unroll
                for (int h = 0; h < H; ++h) {
                    const double2 pv = vp[h * 32];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        x[q][h].x = dds_update_flag<false, ZSAFE>(x[q][h].x, pv.x, tq[q].x, tq[q].y, false, bad);
                        x[q][h].y = dds_update_flag<false, ZSAFE>(x[q][h].y, pv.y, tq[q].x, tq[q].y, false, bad);
                    }
                }
            } else {
#pragma unroll
                for (int h = 0; h < H; ++h) {
                    const double2 pv = vp[h * 32];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const bool inc = (flags >> q) & 1u;
                        x[q][h].x = dds_update_flag<true, ZSAFE>(x[q][h].x, pv.x, tq[q].x, tq[q].y, inc, bad);
                        x[q][h].y = dds_update_flag<true, ZSAFE>(x[q][h].y, pv.y, tq[q].x, tq[q].y, inc, bad);
                    }
                }
            }
            t += L;
            vp += H * 32;
        }
        // the last three updates of rows r+1..r+3 use the rows of the block itself
#pragma unroll
        for (int j = 0; j < 3; ++j) {             // pivot row r + j, table row r + j
#pragma unroll
            for (int q = j + 1; q < 4; ++q) {
                const double2 tp = t[q];
                const bool inc = (fm[r + j] >> (r + q)) & 1ull;
#pragma unroll
                for (int h = 0; h < H; ++h) {
                    x[q][h].x = dds_update_flag<true, ZSAFE>(x[q][h].x, x[j][h].x, tp.x, tp.y, inc, bad);
                    x[q][h].y = dds_update_flag<true, ZSAFE>(x[q][h].y, x[j][h].y, tp.x, tp.y, inc, bad);
                }
            }
            t += L;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int h = 0; h < H; ++h) bl[((r + q) * H + h) * 32] = x[q][h];
    }
#ifndef MPX_DDS_PROBE_NO_REDO     // probe build: sweep without the true-division fallback
    // (a lane beyond column k swept zeros, which produce -0: its flag means nothing.  Lanes with
    // only one of two pairs inside do not occur: the launcher gives the columns behind the last
    // multiple of 128 to the one-pair instantiation.)
    if (bad && active) dds_redo_columns<H>(bl, vals, ld, col0, active, rows, len, gcol, dy2, fm, L);
#endif
    // Write back rows 1..len-1 (element 0 of a fiber never changes), again one bulk copy per row
    fence_proxy_async_smem();
    __syncwarp();
    if (1 + lane < len)
        bulk_s2g(vals + (int64_t)row_out * ld + cw0, bw + (size_t)(1 + lane) * H * 32, row_bytes);
    for (int rr = 33 + lane; rr < len; rr += 32)
        bulk_s2g(vals + (int64_t)rows[rr] * ld + cw0, bw + (size_t)rr * H * 32, row_bytes);
    bulk_commit_and_wait_read();          // the buffer must outlive the copies that read it
}

#ifdef MPX_TUNING   // measured and dropped: built for the lab bench only
// ---- register-resident fibers (all fibers of the dimension at most MPX_DDS_REG_LEN long) -------
// One warp per fiber, lane = value column, one pass per dimension like the kernel above, but the
// fiber never touches shared memory: a lane keeps its column of the fiber in registers, the code
// for every fiber length is fully unrolled, and the (divisor, reciprocal) pairs arrive as a
// kernel PARAMETER -- with static indices every table entry is an immediate constant-bank
// operand of the DADD/DMUL/DFMA that uses it, so an update issues its four FP64 instructions and
// the range test, nothing else.  (On this GPU every shared-memory load costs the FP64 pipe about
// as much as two DFMAs -- tools/lds_dfma_probe.cu -- and the kernel above needs 1.25 of them per
// update.)  Element r receives its updates i = 1..r in order with the final value of element
// i-1, as in the reference; for a fixed i the updates of r = i..len-1 are independent chains.
struct DdsPairs {
    double2 t[MPX_DDS_REG_LEN * (MPX_DDS_REG_LEN - 1) / 2];    // (l, r) at [r (r - 1) / 2 + l]
};

// The rare cases the fast division cannot take (numerator zero-but-not-plus-zero, subnormal,
// huge, inf, nan; a divisor whose reciprocal was not tabulated) do not branch inside the unrolled
// code: every update computes the fast quotient and ORs "this one was not valid" into one sticky
// predicate; a lane whose predicate is set at the end throws its fiber away and redoes it here
// with true divisions, from the values still unchanged in global memory.
__device__ __noinline__ void dds_fiber_slow(double *__restrict__ vals, int64_t ld, int col,
                                            const int32_t *__restrict__ rows, int len,
                                            const double2 *__restrict__ tab)
{
    for (int r = 1; r < len; ++r) {
        double *pr = vals + (int64_t)rows[r] * ld + col;
        double x = *pr;
        for (int i = 1; i <= r; ++i) {
            // rows below r already hold their final values (written by this loop)
            const double pv = vals[(int64_t)rows[i - 1] * ld + col];
            x = __ddiv_rn(__dsub_rn(x, pv), tab[r * (r - 1) / 2 + (i - 1)].x);
        }
        *pr = x;
    }
}

template <bool CHECK_Y>
__device__ __forceinline__ double dds_update_const(double x, double pv, double d, double y, bool &bad)
{
    const double a = __dsub_rn(x, pv);
    const double q = __dmul_rn(a, y);
    const double r = __fma_rn(-d, q, a);
    const double q2 = __fma_rn(r, y, q);
    const int hi = __double2hiint(a);
    // valid when 2^-900 <= |a| < 2^900, or a is +0 (x == pv: the sequence returns the correctly
    // signed zero), and the divisor qualified (its tabulated reciprocal is not 0)
    const bool in_range = ((unsigned)(hi & 0x7ff00000) - (123u << 20)) < (1800u << 20);
    const bool plus_zero = (hi | __double2loint(a)) == 0;
    bad = bad || !(in_range || plus_zero);
    if constexpr (CHECK_Y) bad = bad || __double2hiint(y) == 0;
    return q2;
}

template <int LEN, bool CHECK_Y>
__device__ __forceinline__ void dds_fiber_in_registers(double *__restrict__ vals, int64_t ld, int col,
                                                       const int32_t *__restrict__ rows,
                                                       const DdsPairs &tab)
{
    double *p[LEN];
    double v[LEN];
#pragma unroll
    for (int r = 0; r < LEN; ++r) p[r] = vals + (int64_t)rows[r] * ld + col;
#pragma unroll
    for (int r = 0; r < LEN; ++r) v[r] = *p[r];
    bool bad = false;
#pragma unroll
    for (int i = 1; i < LEN; ++i) {
#pragma unroll
        for (int r = i; r < LEN; ++r) {
#ifdef MPX_DDS_PROBE_ONE_PAIR
            const double2 t = tab.t[0];
#else
            const double2 t = tab.t[r * (r - 1) / 2 + (i - 1)];
#endif
            v[r] = dds_update_const<CHECK_Y>(v[r], v[i - 1], t.x, t.y, bad);
        }
    }
    if (bad) {
        dds_fiber_slow(vals, ld, col, rows, LEN, tab.t);
        return;
    }
#pragma unroll
    for (int r = 1; r < LEN; ++r) *p[r] = v[r];          // element 0 of a fiber never changes
}

template <bool CHECK_Y, int FW>
__global__ void __launch_bounds__(FW * 32)
dds_fiber_reg_kernel(double *__restrict__ vals, int64_t ld, int k, const int32_t *__restrict__ woff,
                     const int32_t *__restrict__ wrows, int n_work,
                     const __grid_constant__ DdsPairs tab)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = blockIdx.x * FW + warp;
    const int col = blockIdx.y * MPX_DDS_COLS + lane;
    if (f >= n_work || col >= k) return;
    const int o = woff[f];
    const int len = woff[f + 1] - o;
    const int32_t *rows = wrows + o;
    switch (len) {
#define MPX_DDS_LEN(L) case L: dds_fiber_in_registers<L, CHECK_Y>(vals, ld, col, rows, tab); break;
        MPX_DDS_LEN(2) MPX_DDS_LEN(3) MPX_DDS_LEN(4) MPX_DDS_LEN(5) MPX_DDS_LEN(6) MPX_DDS_LEN(7)
        MPX_DDS_LEN(8) MPX_DDS_LEN(9) MPX_DDS_LEN(10) MPX_DDS_LEN(11) MPX_DDS_LEN(12)
        MPX_DDS_LEN(13) MPX_DDS_LEN(14) MPX_DDS_LEN(15) MPX_DDS_LEN(16)
#undef MPX_DDS_LEN
    default: break;
    }
}

#endif  // MPX_TUNING

}  // namespace

// several columns per lane: k, ld and the base address must allow 16-byte accesses, fibers <= 40
template <int H, int FW = 4>
static int launch_dds_cols_h(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream,
                             bool *handled, bool zsafe = false)
{
    int smem_max = 0;
    MPX_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, pl->device));
    *handled = false;
    auto smem_of = [&](const DdsDim &dd) {
        return ((size_t)2 * dd.max_len * dd.max_len + (size_t)FW * dd.max_len * 64 * H + FW + dd.max_len) * sizeof(double);
    };
    for (int d = 0; d < pl->m; ++d) {
        const DdsDim &dd = pl->dds[d];
        if (dd.n_work == 0) continue;
        if (dd.max_len > 40 || !dd.cols_ok || smem_of(dd) > (size_t)smem_max) return MPX_OK;
    }
    auto fn = zsafe ? dds_fiber_cols_kernel<FW, H, true> : dds_fiber_cols_kernel<FW, H, false>;
    MPX_CUDA(mpx_allow_max_smem(fn));
    // (Running all passes over one L2-sized group of columns before the next group was measured:
    // 3 % slower at cfg4 -- the passes are not bound by HBM alone, and smaller launches have
    // longer tails.)
    for (int d = pl->m - 1; d >= 0; --d) {
        const DdsDim &dd = pl->dds[d];
        if (dd.n_work == 0) continue;
        dim3 grid((unsigned)((dd.n_work + FW - 1) / FW), (unsigned)((k + 64 * H - 1) / (64 * H)), 1);
        fn<<<grid, FW * 32, smem_of(dd), stream>>>(vals, ld, k, dd.d_woff, dd.d_wpad, dd.n_work,
                                                   dd.d_dy2, dd.d_fmask, dd.d_gcol, dd.max_len);
        MPX_CUDA(cudaGetLastError());
    }
    *handled = true;
    return MPX_OK;
}

static int launch_dds_cols(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream,
                           bool *handled, bool zsafe)
{
    *handled = false;
    if ((k & 1) || (ld & 1) || (reinterpret_cast<uintptr_t>(vals) & 15) || k < 34) return MPX_OK;
#ifdef MPX_TUNING
    static const int fw = getenv("MPX_DDS_COLS_FW") ? atoi(getenv("MPX_DDS_COLS_FW")) : 4;
    static const int hh = getenv("MPX_DDS_COLS_H") ? atoi(getenv("MPX_DDS_COLS_H")) : 0;
    if (hh == 1 || (hh == 0 && k < 128)) {
        if (fw == 2) return launch_dds_cols_h<1, 2>(pl, vals, k, ld, stream, handled);
        if (fw == 8) return launch_dds_cols_h<1, 8>(pl, vals, k, ld, stream, handled);
        return launch_dds_cols_h<1, 4>(pl, vals, k, ld, stream, handled);
    }
    if (fw == 2) return launch_dds_cols_h<2, 2>(pl, vals, k, ld, stream, handled);
    if (fw == 8) return launch_dds_cols_h<2, 8>(pl, vals, k, ld, stream, handled);
    return launch_dds_cols_h<2, 4>(pl, vals, k, ld, stream, handled);
#else
    if (k < 128) return launch_dds_cols_h<1>(pl, vals, k, ld, stream, handled, zsafe);
    // two pairs per lane over the multiples of 128 columns, one pair per lane over the rest
    const int k2 = k & ~127;
    int rc = launch_dds_cols_h<2>(pl, vals, k2, ld, stream, handled, zsafe);
    if (rc != MPX_OK || !*handled || k2 == k) return rc;
    rc = launch_dds_cols_h<1>(pl, vals + k2, k - k2, ld, stream, handled, zsafe);
    if (rc == MPX_OK && !*handled)
        return mpx_fail(MPX_ERR_CUDA, "dds: the one-pair kernel refused what the two-pair kernel took");
    return rc;
#endif
}

#ifdef MPX_TUNING
// register-resident fibers: every dimension with work must have max_len <= MPX_DDS_REG_LEN
static int launch_dds_registers(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream,
                                bool *handled)
{
    *handled = false;
    for (int d = 0; d < pl->m; ++d)
        if (pl->dds[d].n_work > 0 && pl->dds[d].h_pairs.empty()) return MPX_OK;
    constexpr int FW = 4;
    for (int d = pl->m - 1; d >= 0; --d) {
        const DdsDim &dd = pl->dds[d];
        if (dd.n_work == 0) continue;
        DdsPairs tab;
        static_assert(sizeof(tab) == sizeof(double) * MPX_DDS_REG_LEN * (MPX_DDS_REG_LEN - 1), "table size");
        memcpy(&tab, dd.h_pairs.data(), sizeof(tab));
        dim3 grid((unsigned)((dd.n_work + FW - 1) / FW), (unsigned)((k + MPX_DDS_COLS - 1) / MPX_DDS_COLS), 1);
        if (dd.all_fast)
            dds_fiber_reg_kernel<false, FW><<<grid, FW * 32, 0, stream>>>(vals, ld, k, dd.d_woff, dd.d_wrows, dd.n_work, tab);
        else
            dds_fiber_reg_kernel<true, FW><<<grid, FW * 32, 0, stream>>>(vals, ld, k, dd.d_woff, dd.d_wrows, dd.n_work, tab);
        MPX_CUDA(cudaGetLastError());
    }
    *handled = true;
    return MPX_OK;
}

#endif  // MPX_TUNING

template <int FW>
static int launch_dds_fibers_fw(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream,
                                bool *handled)
{
#ifdef MPX_TUNING
    static const bool force_div = getenv("MPX_DDS_TRUE_DIV") != nullptr;
#else
    constexpr bool force_div = false;
#endif
    int smem_max = 0;
    MPX_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, pl->device));
    *handled = false;
    auto smem_of = [&](const DdsDim &dd, bool fast) {
        return ((fast ? (size_t)2 * dd.max_len * dd.max_len : (size_t)((dd.max_len + 1) & ~1)) +
                (size_t)FW * dd.max_len * MPX_DDS_COLS) * sizeof(double);
    };
    // every dimension must fit before anything is launched (no half-transformed arrays)
    for (int d = 0; d < pl->m; ++d) {
        const DdsDim &dd = pl->dds[d];
        if (dd.n_work == 0) continue;
        const bool fast = dd.max_len <= 40 && !force_div;
        if (smem_of(dd, fast) > (size_t)smem_max || dd.max_len > DDS_MAX_FIBER) return MPX_OK;
    }
    for (int d = pl->m - 1; d >= 0; --d) {
        const DdsDim &dd = pl->dds[d];
        if (dd.n_work == 0) continue;
        const bool fast = dd.max_len <= 40 && !force_div;
        const size_t smem = smem_of(dd, fast);
        auto fn = fast ? dds_fiber_kernel<true, FW> : dds_fiber_kernel<false, FW>;
        MPX_CUDA(mpx_allow_max_smem(fn));
        dim3 grid((unsigned)((dd.n_work + FW - 1) / FW), (unsigned)((k + MPX_DDS_COLS - 1) / MPX_DDS_COLS), 1);
        fn<<<grid, FW * 32, smem, stream>>>(vals, ld, k, dd.d_woff, dd.d_wrows, dd.n_work, dd.d_dy,
                                            dd.d_gcol, dd.max_len);
        MPX_CUDA(cudaGetLastError());
    }
    *handled = true;
    return MPX_OK;
}

static int launch_dds_fibers(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream,
                             bool *handled)
{
    // warps per CTA: short fibers leave room for many small CTAs (more resident warps, measured
    // 7 % faster at cfg4); long fibers want fewer, larger buffers per CTA
    int longest = 0;
    for (int d = 0; d < pl->m; ++d) longest = std::max(longest, pl->dds[d].max_len);
#ifdef MPX_TUNING
    static const int forced = getenv("MPX_DDS_FIBER_WARPS") ? atoi(getenv("MPX_DDS_FIBER_WARPS")) : 0;
#else
    constexpr int forced = 0;
#endif
    const int fw = forced ? forced : (longest <= 16 || longest > 64) ? 4 : 8;
    if (fw == 4) return launch_dds_fibers_fw<4>(pl, vals, k, ld, stream, handled);
    if (fw == 16) return launch_dds_fibers_fw<16>(pl, vals, k, ld, stream, handled);
    return launch_dds_fibers_fw<8>(pl, vals, k, ld, stream, handled);
}

#ifdef MPX_TUNING
static int launch_dds_per_dim(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream)
{
    for (int d = pl->m - 1; d >= 0; --d) {
        const DdsDim &dd = pl->dds[d];
        if (dd.max_len < 2) continue;
        const int64_t threads = (int64_t)dd.n_fibers * k;
        const int tb = 256;
        const unsigned blocks = (unsigned)((threads + tb - 1) / tb);
        if (dd.max_len <= 9)
            dds_pass_kernel<9><<<blocks, tb, 0, stream>>>(vals, ld, k, dd.d_off, dd.d_rows,
                                                          dd.n_fibers, dd.d_diff, dd.max_len);
        else if (dd.max_len <= 17)
            dds_pass_kernel<17><<<blocks, tb, 0, stream>>>(vals, ld, k, dd.d_off, dd.d_rows,
                                                           dd.n_fibers, dd.d_diff, dd.max_len);
        else if (dd.max_len <= 33)
            dds_pass_kernel<33><<<blocks, tb, 0, stream>>>(vals, ld, k, dd.d_off, dd.d_rows,
                                                           dd.n_fibers, dd.d_diff, dd.max_len);
        else
            dds_pass_generic_kernel<<<blocks, tb, 0, stream>>>(vals, ld, k, dd.d_off, dd.d_rows,
                                                               dd.n_fibers, dd.d_diff, dd.max_len);
        MPX_CUDA(cudaGetLastError());
    }
    return MPX_OK;
}
#endif

// Test hook (not in the public header): 1 routes mpx_dds through the fused slice kernel, which
// otherwise only runs for index sets whose fibers do not fit the one-warp-per-fiber kernels;
// 2 through the shared-memory fiber kernel, which otherwise only takes fibers longer than 16.
static int mpx_test_dds_group = 0;
extern "C" void mpx_debug_set_dds_group(int on) { mpx_test_dds_group = on; }

int mpx_launch_dds(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream, int flags)
{
    if (!pl->complete)
        return mpx_fail(MPX_ERR_INCOMPLETE,
                        "trying to use the divided difference scheme with incomplete multi "
                        "indices, but DDS only works for complete multi indices (without 'holes').");
    // default: one warp per fiber, one pass per dimension; the fused slice kernel takes over
    // when a fiber buffer does not fit shared memory (and, in a -DMPX_TUNING build, on request)
#ifdef MPX_TUNING
    static const bool per_dim = getenv("MPX_DDS_PER_DIM") != nullptr;   // old one-pass-per-dimension path
    if (per_dim) return launch_dds_per_dim(pl, vals, k, ld, stream);
    static const char *mode = getenv("MPX_DDS_MODE");          // "fiber" | "group"
#else
    const char *mode = mpx_test_dds_group == 1 ? "group" : mpx_test_dds_group == 2 ? "fiber"
                       : mpx_test_dds_group == 3 ? "registers" : nullptr;
#endif
    if (!mode || strcmp(mode, "group") != 0) {
        bool handled = false;
        int rc = MPX_OK;
        if (!mode) {
            rc = launch_dds_cols(pl, vals, k, ld, stream, &handled, (flags & MPX_DDS_MANY_ZEROS) != 0);
            if (rc != MPX_OK || handled) return rc;
        }
#ifdef MPX_TUNING
        if (mode && strcmp(mode, "registers") == 0) {
            rc = launch_dds_registers(pl, vals, k, ld, stream, &handled);
            if (rc != MPX_OK || handled) return rc;
        }
#endif
        rc = launch_dds_fibers(pl, vals, k, ld, stream, &handled);
        if (rc != MPX_OK || handled) return rc;
    }
    for (const DdsGroup &grp : pl->dds_groups) {
        DdsGroupArgs a;
        a.slice_off = grp.d_slice_off;
        a.slice_rows = grp.d_slice_rows;
        a.n_dims = (int)grp.dims.size();
        a.max_rows = grp.max_rows;
        size_t diff_doubles = 0, g_doubles = 0, meta_words = 0;
#ifdef MPX_TUNING
        static const bool force_div = getenv("MPX_DDS_TRUE_DIV") != nullptr;
#else
        constexpr bool force_div = false;
#endif
        bool fast = pl->dds_fast_div && !force_div;
        bool any = false;
        for (int t = 0; t < a.n_dims; ++t) {
            const DdsGroupDim &gd = grp.dims[t];
            a.dim[t] = {gd.d_slice_fib0, gd.d_fib_off, gd.d_fib_rows, gd.d_diff, gd.d_recip, gd.d_g, gd.L};
            a.max_fibers[t] = gd.max_fibers;
            fast = fast && gd.L <= 40;          // L x L tables of diff and recip must stay small
            diff_doubles += (size_t)2 * gd.L * gd.L;
            g_doubles += (size_t)gd.L;
            meta_words += (size_t)gd.max_fibers + 1 + grp.max_rows;
            any = any || gd.L >= 2;
            if (gd.L > DDS_MAX_FIBER) return mpx_fail(MPX_ERR_UNSUPPORTED, "fiber longer than 256");
        }
        if (!any) continue;      // every fiber of this group has a single element
        const size_t smem = ((size_t)grp.max_rows * MPX_DDS_COLS + (fast ? diff_doubles : g_doubles)) * sizeof(double) +
                            meta_words * sizeof(uint32_t);
        int smem_max = 0;
        MPX_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, pl->device));
        if (smem > (size_t)smem_max) return mpx_fail(MPX_ERR_UNSUPPORTED, "dds slice does not fit shared memory");
        auto fn = fast ? dds_group_kernel<true> : dds_group_kernel<false>;
        MPX_CUDA(mpx_allow_max_smem(fn));
        dim3 grid((unsigned)grp.n_slices, (unsigned)((k + MPX_DDS_COLS - 1) / MPX_DDS_COLS), 1);
        fn<<<grid, DDS_WARPS * 32, smem, stream>>>(vals, ld, k, a);
        MPX_CUDA(cudaGetLastError());
    }
    return MPX_OK;
}
