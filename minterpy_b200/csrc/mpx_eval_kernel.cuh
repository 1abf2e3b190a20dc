// Tree-form evaluation kernel (K1) and its launcher, shared by mpx_eval.cu (plain evaluation) and
// mpx_eval_deriv.cu (partial derivatives).  See mpx_eval.cu for the description.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <type_traits>

#include "mpx_common.h"
#include "mpx_ptx.h"

namespace {

struct EvalParams {
    const double *xx;
    int64_t ld_xx;
    int64_t P;
    double *out;
    int64_t ld_out;
    const uint32_t *desc;   // n_chunks * MPX_DESC_WORDS
    const double *coef;     // k * n_chunks * MPX_COEF_CAP (packed per call)
    const double *gp;       // m x gp_rows, dimension-major
    int n_chunks;
    int stages;             // ring depth actually used (<= MPX_STAGES)
    int m;
    int gp_rows;
    int max_exp[MPX_MAX_M];
    int start_exp[MPX_MAX_M];
    // derivative evaluation only (DERIV kernels): orders per dimension, and the tables of the
    // dimensions >= 2 -- entry e of dimension 2+u for point p at dtab[(u * gp_rows + e) * P + p]
    int order[MPX_MAX_M];
    const double *dtab;
};

constexpr int STAGE_BYTES = MPX_DESC_WORDS * 4 + MPX_COEF_CAP * 8;

// pack kernel: coefficient image [q][chunk][slot] from the caller's (N x k) matrix
__global__ void pack_coeffs_kernel(const double *__restrict__ coeffs, int64_t ld_c, int k,
                                   const int32_t *__restrict__ slot_src, int64_t n_slots,
                                   double *__restrict__ img)
{
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_slots * k) return;
    int q = (int)(t % k);
    int64_t s = t / k;
    int32_t src = slot_src[s];
    img[(int64_t)q * n_slots + s] = (src >= 0) ? coeffs[(int64_t)src * ld_c + q] : 0.0;
}

// T1 (dimension-1 table) in shared memory.  R even: [e][r/2][thread][r%2], so one LDS.128
// fetches the values of two points (lane stride 16 B, conflict free); R odd: [e][r][thread].
template <int R, int NT>
struct T1Layout {
    static constexpr bool PAIRED = (R % 2 == 0);
    static constexpr uint32_t ROW_BYTES = (uint32_t)R * NT * 8u;
    static __device__ __forceinline__ uint32_t index(int e, int r, int tid)
    {
        if constexpr (PAIRED) return ((uint32_t)(e * (R / 2) + r / 2) * NT + tid) * 2 + (r & 1);
        else return (uint32_t)(e * R + r) * NT + tid;
    }
    static __device__ __forceinline__ uint32_t lane_base(uint32_t t1s_addr, int tid)
    {
        return t1s_addr + (uint32_t)tid * (PAIRED ? 16u : 8u);
    }
    static __device__ __forceinline__ void load(double (&v)[R], uint32_t row_addr)
    {
        if constexpr (PAIRED) {
#pragma unroll
            for (int h = 0; h < R / 2; ++h) {
                const double2 t = lds128(row_addr + (uint32_t)h * NT * 16u);
                v[2 * h] = t.x;
                v[2 * h + 1] = t.y;
            }
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) v[r] = lds64(row_addr + (uint32_t)r * NT * 8u);
        }
    }
};

// `cnt` consecutive leaves of K coefficient pairs each (dimension-1 exponent counting up).
// K is a compile-time constant, so the body is straight-line code: K broadcast loads of
// coefficient pairs, the T1 values of the R points, 2K FMAs per point into A independent
// accumulators (A = 2: odd and even monomials apart), then one FMA per point into the
// level-1 sum.  Lowest monomial first.
template <int L0P, int R, int A, int NT, int K, bool DERIV>
__device__ __forceinline__ void run_leaves(double (&acc1)[R], const double (&T0)[R][L0P],
                                           uint32_t &cp, uint32_t &t1a, int cnt)
{
#pragma unroll 1
    do {
        double2 cf[K];
#pragma unroll
        for (int i = 0; i < K; ++i) cf[i] = lds128(cp + 16 * i);
        double t1v[R];
        T1Layout<R, NT>::load(t1v, t1a);
        // T0[.][0] is 1: the coefficient of exponent 0 enters as the addend of the first FMA,
        // so a leaf of 2K monomials costs 2K - 1 FP64 operations (plus the level-1 FMA)
        double a[R][A];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            // (a differentiated dimension-0 table does not start with 1: DERIV multiplies)
            const double c0 = DERIV ? cf[0].x * T0[r][0] : cf[0].x;
            if constexpr (A == 2) {
                a[r][1] = cf[0].y * T0[r][1];
                a[r][0] = c0;
            } else {
                a[r][0] = fma(cf[0].y, T0[r][1], c0);
            }
        }
#pragma unroll
        for (int i = 1; i < K; ++i) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                a[r][0] = fma(cf[i].x, T0[r][2 * i], a[r][0]);
                a[r][A - 1] = fma(cf[i].y, T0[r][2 * i + 1], a[r][A - 1]);
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if constexpr (A == 2) acc1[r] = fma(t1v[r], a[r][0] + a[r][1], acc1[r]);
            else acc1[r] = fma(t1v[r], a[r][0], acc1[r]);
        }
        cp += 16 * K;
        t1a += T1Layout<R, NT>::ROW_BYTES;
    } while (--cnt);
}

// walk the lengths of one record downwards: K = L0P/2, ..., 1
template <int L0P, int R, int A, int NT, int K, bool DERIV>
struct LengthChain {
    static __device__ __forceinline__ void go(double (&acc1)[R], const double (&T0)[R][L0P],
                                              uint32_t &cp, uint32_t &t1a,
                                              const uint32_t (&counts)[(L0P / 2 + 7) / 8])
    {
        // entering a count word (8 lengths) that is all zero: skip its lengths in one test
        constexpr bool word_top = ((K - 1) % 8 == 7) || (K == L0P / 2);
        constexpr int below = ((K - 1) / 8) * 8;         // first length of the next lower word
        if (word_top && counts[(K - 1) / 8] == 0u) {
            if constexpr (below >= 1) LengthChain<L0P, R, A, NT, (below >= 1 ? below : 1), DERIV>::go(acc1, T0, cp, t1a, counts);
            return;
        }
        const int cnt = (int)((counts[(K - 1) / 8] >> (4 * ((K - 1) % 8))) & 15u);
        if (cnt) run_leaves<L0P, R, A, NT, K, DERIV>(acc1, T0, cp, t1a, cnt);
        if constexpr (K > 1) LengthChain<L0P, R, A, NT, K - 1, DERIV>::go(acc1, T0, cp, t1a, counts);
    }
};

// shared-memory carve-up (dynamic):
//   [0, 128)                     mbarriers: full[], empty[]
//   [128, 128 + S*STAGE_BYTES)   ring: per stage desc words then coefficients
//   then gp (m*gp_rows doubles), T1 ((max_exp1+1)*R*NT doubles),
//   running product / partial sum of dimensions 3.. (2 * (U-1) * R * NT doubles; dimension 2
//   keeps its pair in registers)
// UP = MpxUniformProgram: the record words come from a kernel parameter (constant bank) instead
// of the shared-memory ring and records have a fixed size, so everything the control flow of the
// leaf loops depends on -- counts, fold flags, loop bounds -- is provably warp-uniform: ptxas
// keeps it in uniform registers and branches uniformly (BRA.U, no BSSY/BSYNC brackets).  That
// only happens if the warp does not spin on the ring's mbarrier with all 32 lanes (a per-thread
// loop exit makes ptxas treat what follows as divergent): lane 0 waits, the warp meets.
struct NoUniformProgram {};
template <int L0P, int R, int A, int NT, bool DERIV, class UP>
__device__ __forceinline__ void eval_tree_body(const EvalParams &p, const UP &up)
{
    constexpr bool UDESC = !std::is_same<UP, NoUniformProgram>::value;
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int NCW = (L0P / 2 + 7) / 8;
    using T1L = T1Layout<R, NT>;
    const int tid = threadIdx.x;
    const int m = p.m;
    const int U = m > 2 ? m - 2 : 0;
    const int S = p.stages;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + MPX_STAGES;
    unsigned char *ring = smem + 128;
    double *gps = reinterpret_cast<double *>(ring + S * STAGE_BYTES);
    double *T1s = gps + ((m * p.gp_rows + 1) & ~1);
    const int me1 = (m >= 2) ? p.max_exp[1] : 0;
    double *ups = T1s + (size_t)(me1 + 1) * R * NT;   // [2][U-1][R][NT]: running product, sum
    const size_t up_stride = (size_t)(U > 1 ? U - 1 : 0) * R * NT;   // dimensions 3.. only
    // coordinates of dimensions 2.. of this thread's points, [U][R][NT]: the folds read them
    // with one address computation instead of rebuilding R global pointers per record
    double *xs = ups + 2 * up_stride + tid;

    const int q = blockIdx.y;                                   // polynomial
    const double *coef_img = p.coef + (size_t)q * p.n_chunks * MPX_COEF_CAP;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NT);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < m * p.gp_rows; i += NT) gps[i] = p.gp[i];
    __syncthreads();
    // Producer = thread 0, never blocking unless it must: it tops the ring up at every chunk
    // boundary with whatever stages all threads have released by then.
    int next_load = 0;
    auto issue_load = [&](int cn) {
        const int sp = cn % S;
        unsigned char *dst = ring + sp * STAGE_BYTES;
        mbar_arrive_expect_tx(&full[sp], STAGE_BYTES);
        bulk_g2s(dst, p.desc + (size_t)cn * MPX_DESC_WORDS, MPX_DESC_WORDS * 4, &full[sp]);
        bulk_g2s(dst + MPX_DESC_WORDS * 4, coef_img + (size_t)cn * MPX_COEF_CAP,
                 MPX_COEF_CAP * 8, &full[sp]);
    };
    if (tid == 0) {
        const int pre = p.n_chunks < S ? p.n_chunks : S;
        for (; next_load < pre; ++next_load) issue_load(next_load);
    } else {
        next_load = p.n_chunks;      // only thread 0 ever refills
    }
    // "thread 0 still has chunks to request": one predicate for the polls in the record loop
    // instead of a parameter load and a compare per record
    bool refilling = next_load < p.n_chunks;

    // ---- per-point tables -------------------------------------------------------------
    const int64_t tile0 = (int64_t)blockIdx.x * (R * NT);
    double T0[R][L0P];
    double acc1[R];
    double t2[R], a2[R];          // running product and partial sum of dimension 2
    // DERIV: tables of the dimensions >= 2 in global memory; this thread's points in them
    const double *dt[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int64_t pt = tile0 + (int64_t)r * NT + tid;
        if (pt >= p.P) pt = p.P - 1;               // tail threads recompute the last point
        dt[r] = DERIV ? p.dtab + pt : nullptr;
        const double *xp = p.xx + pt * p.ld_xx;
        const double x0 = xp[0];
        const int me0 = p.max_exp[0];
        T0[r][0] = 1.0;
#pragma unroll
        for (int j = 0; j < L0P - 1; ++j)
            T0[r][j + 1] = (j < me0) ? T0[r][j] * (x0 - gps[j]) : 0.0;
        if constexpr (DERIV) {
            // d^o/dx^o of the table entries by the product rule on T[q+1] = (x - g_q) T[q]:
            // F_j[q+1] = (x - g_q) F_j[q] + F_{j-1}[q], one pass per order, in place
            // (csrc/mpx_eval_mma.cu: derivative_table), then the factor o!
            const int o0 = p.order[0];
            double fact = 1.0;
#pragma unroll 1
            for (int pass = 1; pass <= o0; ++pass) {
                double fj = 0.0, old = T0[r][0];
                T0[r][0] = 0.0;
#pragma unroll
                for (int j = 0; j < L0P - 1; ++j) {
                    const double old_next = T0[r][j + 1];
                    fj = (j < me0) ? fma(x0 - gps[j], fj, old) : 0.0;
                    T0[r][j + 1] = fj;
                    old = old_next;
                }
                fact *= (double)pass;
            }
#pragma unroll
            for (int j = 0; j < L0P; ++j) T0[r][j] *= fact;
        }
        T1s[T1L::index(0, r, tid)] = 1.0;
        if (m >= 2) {
            const double x1 = xp[1];
            double prod = 1.0;
#pragma unroll 1
            for (int j = 0; j < me1; ++j) {
                prod *= x1 - gps[p.gp_rows + j];
                T1s[T1L::index(j + 1, r, tid)] = prod;
            }
            if constexpr (DERIV) {
                const int o1 = p.order[1];
                double fact = 1.0;
#pragma unroll 1
                for (int pass = 1; pass <= o1; ++pass) {
                    double fj = 0.0, old = T1s[T1L::index(0, r, tid)];
                    T1s[T1L::index(0, r, tid)] = 0.0;
#pragma unroll 1
                    for (int j = 0; j < me1; ++j) {
                        const double old_next = T1s[T1L::index(j + 1, r, tid)];
                        fj = fma(x1 - gps[p.gp_rows + j], fj, old);
                        T1s[T1L::index(j + 1, r, tid)] = fj;
                        old = old_next;
                    }
                    fact *= (double)pass;
                }
                if (o1 > 0)
#pragma unroll 1
                    for (int j = 0; j <= me1; ++j) T1s[T1L::index(j, r, tid)] *= fact;
            }
        }
#pragma unroll 1
        for (int u = 0; u < U; ++u) {
            const double xu = xp[2 + u];
            xs[(size_t)(u * R + r) * NT] = xu;
            double t = 1.0;
            if constexpr (DERIV) {
                t = dt[r][(size_t)(u * p.gp_rows + p.start_exp[2 + u]) * p.P];
            } else {
#pragma unroll 1
                for (int j = 0; j < p.start_exp[2 + u]; ++j) t *= xu - gps[(2 + u) * p.gp_rows + j];
            }
            if (u == 0) {
                t2[r] = t;               // dimension 2 is folded into often: registers
            } else {
                const size_t o = ((size_t)(u - 1) * R + r) * NT + tid;
                ups[o] = t;
                ups[up_stride + o] = 0.0;
            }
        }
        a2[r] = 0.0;
        acc1[r] = 0.0;
    }
    // (each thread only ever touches its own T1s / ups entries: no barrier needed)
    const uint32_t t1_base = T1L::lane_base(smem_u32(T1s), tid);

    // ---- stream the program -------------------------------------------------------------
    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % S;
        const uint32_t ph = (uint32_t)(c / S) & 1u;
        if (refilling) {
            // thread 0: refill every stage that all threads have released; if chunk c itself
            // has not been requested yet (its stage was busy at every earlier poll), wait.
            for (;;) {
                const int sp = next_load % S;
                const uint32_t php = (uint32_t)(next_load / S - 1) & 1u;
                if (next_load <= c) mbar_wait(&empty[sp], php);
                else if (!mbar_test(&empty[sp], php)) break;
                issue_load(next_load);
                if (++next_load >= p.n_chunks) { refilling = false; break; }
            }
        }
        if constexpr (UDESC) {
            if ((tid & 31) == 0) mbar_wait(&full[s], ph);
            __syncwarp();
        } else {
            mbar_wait(&full[s], ph);
        }
        const uint32_t *dwp = reinterpret_cast<const uint32_t *>(ring + s * STAGE_BYTES);
        uint32_t ubase = 0;
        if constexpr (UDESC) ubase = up.off[c];
        auto dw = [&](int idx) -> uint32_t {
            if constexpr (UDESC) return up.w[ubase + idx];
            else return dwp[idx];
        };
        uint32_t cp = smem_u32(ring + s * STAGE_BYTES + MPX_DESC_WORDS * 4);
        const int iend = MPX_DESC_HDR + (int)dw(0);
        // UDESC: records have a fixed size (no irregular ones, the plan checks), so the word index
        // is a function of a plain loop counter
        const int nrec = (iend - MPX_DESC_HDR) / (1 + NCW);
        for (int rec = 0, i = MPX_DESC_HDR; UDESC ? rec < nrec : i < iend; ++rec) {
            if constexpr (UDESC) i = MPX_DESC_HDR + rec * (1 + NCW);
            if (refilling) {
                // thread 0, between records: hand a released stage straight back to the TMA
                const int sp = next_load % S;
                if (mbar_test(&empty[sp], (uint32_t)(next_load / S - 1) & 1u)) {
                    issue_load(next_load);
                    refilling = ++next_load < p.n_chunks;
                }
            }
            const uint32_t w = dw(i);
            uint32_t counts[NCW];
#pragma unroll
            for (int z = 0; z < NCW; ++z) counts[z] = dw(i + 1 + z);
            i += 1 + NCW;
            const int e1 = w & 255;
            const int nf = (w >> 8) & 15;
            // a fold follows this record: fetch the coordinate of the level that will advance
            // now, so the load is hidden behind the record's FMAs
            double xadv[R];
            // the common fold -- dimension 2 takes the record's sum and advances by one, nothing
            // irregular, not the end of the program -- gets a short path of its own; what it
            // needs is fetched here, behind the record's FMAs
            const bool simple_fold = (w & (0xF00u | MPX_D_IRREG | MPX_D_LAST)) == 0x100u;
            double g2s = 0.0;
            if (nf) {
#pragma unroll
                for (int r = 0; r < R; ++r) xadv[r] = xs[(size_t)((nf - 1) * R + r) * NT];
                if (simple_fold) g2s = gps[2 * p.gp_rows + (int)((w >> 14) & 255u)];
            }
            uint32_t t1a = t1_base + (uint32_t)e1 * T1L::ROW_BYTES;
            LengthChain<L0P, R, A, NT, L0P / 2, DERIV>::go(acc1, T0, cp, t1a, counts);
            if (simple_fold) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    a2[r] = fma(t2[r], acc1[r], a2[r]);
                    if constexpr (DERIV) t2[r] = dt[r][(size_t)(((w >> 14) & 255u) + 1u) * p.P];
                    else t2[r] *= xadv[r] - g2s;
                    acc1[r] = 0.0;
                }
            } else if (nf) {
                // fold finished subtrees upwards: levels 0..nf-1 (dimensions 2..nf+1)
                int adv = 1, cur = (w >> 14) & 255;
                const bool irregular = !UDESC && (w & MPX_D_IRREG) != 0;
                const bool last = (w & MPX_D_LAST) != 0;
                if (irregular) {
                    const uint32_t x = dw(i);
                    cur = x & 255;
                    adv = (int)(x >> 8);
                }
                double low[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    low[r] = acc1[r];
                    acc1[r] = 0.0;
                }
                if (nf == 1) {
                    // the common fold: dimension 2 takes the sum and moves on (registers only)
                    const double g2 = gps[2 * p.gp_rows + cur];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        a2[r] = fma(t2[r], low[r], a2[r]);
                        if constexpr (DERIV) {
                            if (!last) t2[r] = dt[r][(size_t)(cur + adv) * p.P];
                        } else {
                            if (!last) t2[r] *= xadv[r] - g2;
                        }
                    }
                    if (!DERIV && adv > 1 && !last) {
#pragma unroll 1
                        for (int j = 1; j < adv; ++j) {
                            const double gj = gps[2 * p.gp_rows + cur + j];
#pragma unroll
                            for (int r = 0; r < R; ++r) t2[r] *= xadv[r] - gj;
                        }
                    }
                } else {
                    // dimension 2 is complete as well: hand its sum upwards and restart it
                    {
                        const int st = irregular ? (int)dw(i + 1) : 0;
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            low[r] = fma(t2[r], low[r], a2[r]);
                            a2[r] = 0.0;
                            double tn = 1.0;
                            if constexpr (DERIV) {
                                tn = dt[r][(size_t)st * p.P];
                            } else if (st) {
                                const double xu = xs[(size_t)r * NT];
#pragma unroll 1
                                for (int j = 0; j < st; ++j) tn *= xu - gps[2 * p.gp_rows + j];
                            }
                            t2[r] = tn;
                        }
                    }
#pragma unroll 1
                    for (int u = 1; u < nf - 1; ++u) {
                        double *ut = ups + (size_t)(u - 1) * R * NT + tid;
                        double tv[R], av[R];
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            tv[r] = ut[r * NT];
                            av[r] = ut[up_stride + r * NT];
                        }
                        const int st = irregular ? (int)dw(i + 1 + u) : 0;
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            low[r] = fma(tv[r], low[r], av[r]);
                            double tn = 1.0;
                            if constexpr (DERIV) {
                                tn = dt[r][(size_t)(u * p.gp_rows + st) * p.P];
                            } else if (st) {
                                const double xu = xs[(size_t)(u * R + r) * NT];
#pragma unroll 1
                                for (int j = 0; j < st; ++j) tn *= xu - gps[(2 + u) * p.gp_rows + j];
                            }
                            ut[up_stride + r * NT] = 0.0;
                            ut[r * NT] = tn;
                        }
                    }
                    {
                        const int u = nf - 1;
                        double *ut = ups + (size_t)(u - 1) * R * NT + tid;
                        const double g = gps[(2 + u) * p.gp_rows + cur];
                        double tv[R], av[R];
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            tv[r] = ut[r * NT];
                            av[r] = ut[up_stride + r * NT];
                        }
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            av[r] = fma(tv[r], low[r], av[r]);
                            if constexpr (DERIV) {
                                // (past the largest exponent only on the last record: unused)
                                const int ne = cur + adv <= p.max_exp[2 + u] ? cur + adv : cur;
                                tv[r] = dt[r][(size_t)(u * p.gp_rows + ne) * p.P];
                            } else {
                                tv[r] *= xadv[r] - g;
                            }
                        }
                        if (!DERIV && adv > 1) {
#pragma unroll 1
                            for (int j = 1; j < adv; ++j) {
                                const double gj = gps[(2 + u) * p.gp_rows + cur + j];
#pragma unroll
                                for (int r = 0; r < R; ++r) tv[r] *= xadv[r] - gj;
                            }
                        }
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            ut[up_stride + r * NT] = av[r];
                            if (!last) ut[r * NT] = tv[r];
                        }
                    }
                }
                if (!UDESC && irregular) i += nf;
            }
        }
        mbar_arrive(&empty[s]);
    }

    // ---- result ---------------------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t pt = tile0 + (int64_t)r * NT + tid;
        if (pt < p.P) {
            const double res = (U == 0) ? acc1[r] : (U == 1) ? a2[r]
                               : ups[up_stride + ((size_t)(U - 2) * R + r) * NT + tid];
            p.out[pt * p.ld_out + q] = res;
        }
    }
}

template <int L0P, int R, int A, int NT, int MINB, bool DERIV>
__global__ void __launch_bounds__(NT, MINB) eval_tree_kernel(const __grid_constant__ EvalParams p)
{
    eval_tree_body<L0P, R, A, NT, DERIV, NoUniformProgram>(p, NoUniformProgram());
}

template <int L0P, int R, int A, int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB)
eval_tree_kernel_u(const __grid_constant__ EvalParams p, const __grid_constant__ MpxUniformProgram up)
{
    eval_tree_body<L0P, R, A, NT, false, MpxUniformProgram>(p, up);
}

// (derivative evaluation builds the 128- and 64-thread variants only: half the code size)
template <int L0P, int R, int A, int NT, int MINB, bool DERIV>
const void *kfn_main()
{
    if constexpr (DERIV) return nullptr;
    else return reinterpret_cast<const void *>(&eval_tree_kernel<L0P, R, A, NT, MINB, DERIV>);
}

struct Variant {
    int l0p, r, nt;
    const void *fn;
    const void *fn_u;     // the same kernel with the record stream as a kernel parameter, or null
};

template <int L0P, int R, int A, int NT, int MINB, bool DERIV>
const void *kfn_u()
{
    if constexpr (DERIV) return nullptr;
    else return reinterpret_cast<const void *>(&eval_tree_kernel_u<L0P, R, A, NT, MINB>);
}

template <int L0P, int R, int A, int NT, int MINB, bool DERIV>
const void *kfn()
{
    return reinterpret_cast<const void *>(&eval_tree_kernel<L0P, R, A, NT, MINB, DERIV>);
}

size_t eval_smem_bytes(const mpx_plan *pl, int R, int NT, int stages)
{
    const int m = pl->m;
    const int U = m > 2 ? m - 2 : 0;
    const int me1 = m >= 2 ? pl->max_exp[1] : 0;
    size_t doubles = (size_t)((m * pl->gp_rows + 1) & ~1) + (size_t)(me1 + 1) * R * NT +
                     (size_t)2 * (U > 1 ? U - 1 : 0) * R * NT + (size_t)U * R * NT;
    return 128 + (size_t)stages * STAGE_BYTES + doubles * 8;
}

// Variant choice, coefficient packing and launch of the tree kernel.  DERIV = false: plain
// evaluation (mpx_eval.cu); DERIV = true: a partial derivative (mpx_eval_deriv.cu), `orders`
// (m entries, host) and `dtab` (tables of the dimensions >= 2, device) then describe it.
template <bool DERIV>
int tree_launch(mpx_plan *pl, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                cudaStream_t stream, const int *orders, const double *dtab)
{
    const int l0p = mpx_l0p_for(pl->max_exp[0]);
    // kernel variants: (dimension-0 table size, points per thread, threads per CTA); for each
    // table size the first entry whose shared memory fits is taken
    static const Variant variants[] = {
        {10, 4, 256, kfn_main<10, 4, 1, 256, 1, DERIV>(), kfn_u<10, 4, 1, 256, 1, DERIV>()},
        {10, 4, 128, kfn<10, 4, 1, 128, 1, DERIV>(), kfn_u<10, 4, 1, 128, 1, DERIV>()},
        {10, 4, 64, kfn<10, 4, 1, 64, 1, DERIV>(), kfn_u<10, 4, 1, 64, 1, DERIV>()},
        {10, 1, 64, kfn<10, 1, 2, 64, 1, DERIV>(), nullptr},      // last resort: very long dimension-1 tables
        {18, 4, 256, kfn_main<18, 4, 1, 256, 1, DERIV>(), kfn_u<18, 4, 1, 256, 1, DERIV>()},
        {18, 4, 128, kfn<18, 4, 1, 128, 1, DERIV>(), kfn_u<18, 4, 1, 128, 1, DERIV>()},
        {18, 4, 64, kfn<18, 4, 1, 64, 1, DERIV>(), kfn_u<18, 4, 1, 64, 1, DERIV>()},
        {18, 1, 64, kfn<18, 1, 2, 64, 1, DERIV>(), nullptr},
        {26, 3, 256, kfn_main<26, 3, 1, 256, 1, DERIV>(), kfn_u<26, 3, 1, 256, 1, DERIV>()},
        {26, 3, 128, kfn<26, 3, 1, 128, 1, DERIV>(), kfn_u<26, 3, 1, 128, 1, DERIV>()},
        {26, 3, 64, kfn<26, 3, 1, 64, 1, DERIV>(), kfn_u<26, 3, 1, 64, 1, DERIV>()},
        {26, 1, 64, kfn<26, 1, 2, 64, 1, DERIV>(), nullptr},
        {34, 2, 256, kfn_main<34, 2, 2, 256, 1, DERIV>(), kfn_u<34, 2, 2, 256, 1, DERIV>()},
        {34, 2, 128, kfn<34, 2, 2, 128, 1, DERIV>(), kfn_u<34, 2, 2, 128, 1, DERIV>()},
        {34, 2, 64, kfn<34, 2, 2, 64, 1, DERIV>(), kfn_u<34, 2, 2, 64, 1, DERIV>()},
        {34, 1, 64, kfn<34, 1, 2, 64, 1, DERIV>(), nullptr},
        {66, 1, 256, kfn_main<66, 1, 2, 256, 1, DERIV>(), nullptr},
        {66, 1, 128, kfn<66, 1, 2, 128, 1, DERIV>(), nullptr},
        {66, 1, 64, kfn<66, 1, 2, 64, 1, DERIV>(), nullptr},
#ifdef MPX_TUNING
        // measured and dropped; built only with -DMPX_TUNING, MPX_EVAL_VARIANT=<index>
        {18, 4, 256, kfn_main<18, 4, 2, 256, 1, DERIV>(), kfn_u<18, 4, 2, 256, 1, DERIV>()},   // 19: two accumulators per point
        {18, 3, 384, kfn<18, 3, 1, 384, 1, DERIV>(), nullptr},   // 20: 3 points, 12 warps
        {18, 2, 512, kfn<18, 2, 1, 512, 1, DERIV>(), nullptr},   // 21: 2 points, 16 warps in one CTA
#endif
    };
    constexpr int n_standard = 19;
    int smem_max = 0;
    MPX_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, pl->device));
    // The shipped library takes no tuning input from the environment: the knobs below exist only
    // in a -DMPX_TUNING build (lab bench) and in the ring stress test's build (MPX_TEST_HOOKS).
#ifdef MPX_TUNING
    static const int env_variant = getenv("MPX_EVAL_VARIANT") ? atoi(getenv("MPX_EVAL_VARIANT")) : -1;
    static const int env_stages = getenv("MPX_EVAL_STAGES") ? atoi(getenv("MPX_EVAL_STAGES")) : 0;
    static const bool no_pair = getenv("MPX_EVAL_NO_CTA_PAIR") != nullptr;
#else
    constexpr int env_variant = -1;
    const int env_stages = mpx_test_eval_stages;       // 0 unless a test set it
    const bool no_pair = mpx_test_eval_stages != 0;
#endif
    int stages = env_stages > 0 ? env_stages : 4;
    if (stages > MPX_STAGES) stages = MPX_STAGES;
    if (stages < 2) stages = 2;
    const Variant *v = nullptr;
    // Index sets with long leaves (>= 8 monomials per leaf: cfg2, cfg5) run 1-2 % faster as two
    // independent CTAs of 128 threads per SM, each with a two-stage ring, than as one CTA of
    // 256: the CTAs drift apart and fill each other's gaps at leaf and record boundaries.
    // Short-leaf sets lose as much (more records per chunk, the shallow ring runs dry).
    if (env_stages == 0 && !no_pair && pl->prog.n_leaves > 0 && pl->N >= 8 * pl->prog.n_leaves) {
        int sm_total = 0;
        MPX_CUDA(cudaDeviceGetAttribute(&sm_total, cudaDevAttrMaxSharedMemoryPerMultiprocessor, pl->device));
        for (int i = 0; i < n_standard && !v; ++i)
            if (variants[i].fn && variants[i].l0p == l0p && variants[i].nt == 128 && variants[i].r > 1 &&
                2 * (eval_smem_bytes(pl, variants[i].r, 128, 2) + 1024) <= (size_t)sm_total) {
                v = &variants[i];
                stages = 2;
            }
    }
    for (int i = 0; i < n_standard && !v; ++i)
        if (variants[i].fn && variants[i].l0p == l0p &&
            eval_smem_bytes(pl, variants[i].r, variants[i].nt, stages) <= (size_t)smem_max)
            v = &variants[i];
    if (env_variant >= 0 && env_variant < (int)(sizeof(variants) / sizeof(variants[0])) &&
        variants[env_variant].l0p == l0p)
        v = &variants[env_variant];
    if (!v) return mpx_fail(MPX_ERR_UNSUPPORTED, "tables of this index set do not fit shared memory");
    const int nt = v->nt;
    const size_t smem = eval_smem_bytes(pl, v->r, nt, stages);
    if (smem > (size_t)smem_max)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "tables of this index set do not fit shared memory");

    // pack the coefficients into the program image (plan scratch)
    const int64_t n_slots = (int64_t)pl->prog.n_chunks * MPX_COEF_CAP;
    if (coeffs) {
        int rc = mpx_image_acquire(img, (size_t)n_slots * k * sizeof(double), stream);
        if (rc != MPX_OK) return rc;
        const int64_t total = n_slots * k;
        const int tb = 256;
        pack_coeffs_kernel<<<(unsigned)((total + tb - 1) / tb), tb, 0, stream>>>(
            coeffs, ld_c, k, pl->d_slot_src, n_slots, img->ptr);
        MPX_CUDA(cudaGetLastError());
    }
    if (P == 0) return coeffs ? mpx_image_release(img, stream) : MPX_OK;

    EvalParams prm;
    prm.xx = xx; prm.ld_xx = ld_xx; prm.P = P; prm.out = out; prm.ld_out = ld_out;
    prm.desc = pl->d_desc; prm.coef = img->ptr; prm.gp = pl->d_gp;
    prm.n_chunks = pl->prog.n_chunks; prm.stages = stages; prm.m = pl->m; prm.gp_rows = pl->gp_rows;
    for (int i = 0; i < MPX_MAX_M; ++i) {
        prm.max_exp[i] = pl->max_exp[i];
        prm.start_exp[i] = pl->start_exp[i];
        prm.order[i] = (DERIV && orders && i < pl->m) ? orders[i] : 0;
    }
    prm.dtab = dtab;

    MPX_CUDA(mpx_allow_max_smem(v->fn));
    const int64_t per_cta = (int64_t)v->r * nt;
    const int64_t tiles = (P + per_cta - 1) / per_cta;
    if (tiles > 0x7fffffffLL || k > 65535)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points or polynomials for one launch");
    dim3 grid((unsigned)tiles, (unsigned)k, 1);
    if (v->fn_u && !pl->uprog.empty() && mpx_test_eval_uniform != 2) {
        MPX_CUDA(mpx_allow_max_smem(v->fn_u));
        void *args[] = {(void *)&prm, (void *)pl->uprog.data()};
        MPX_CUDA(cudaLaunchKernel(v->fn_u, grid, dim3(nt, 1, 1), args, smem, stream));
        return mpx_image_release(img, stream);
    }
    void *args[] = {(void *)&prm};
    MPX_CUDA(cudaLaunchKernel(v->fn, grid, dim3(nt, 1, 1), args, smem, stream));
    return mpx_image_release(img, stream);
}

}  // namespace
