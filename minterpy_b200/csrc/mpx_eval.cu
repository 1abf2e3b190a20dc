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
#include <cstdio>
#include <cstdlib>

#include "mpx_common.h"

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "LAB_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE;\n\t"
        "bra LAB_WAIT;\n\t"
        "DONE:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

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

__device__ __forceinline__ double2 lds128(uint32_t addr)
{
    // volatile: keeps all coefficient loads of a leaf in front of the length switch, so they
    // are in flight together instead of one exposed shared-memory latency per pair
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}

__device__ __forceinline__ double lds64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

// One run: `cnt` leaves of K coefficient pairs each, dimension-1 exponents counting upwards.
// K is a compile-time constant here, so the body is straight-line code: K broadcast loads of
// coefficient pairs, R loads of T1, 2K FMAs per point into A independent accumulators
// (A = 2: odd and even monomials apart, twice the independent chains), then one FMA per
// point into the level-1 sum.  Highest monomial first, as a Duff-style entry would do.
template <int L0P, int R, int A, int K>
struct Run {
    static __device__ __forceinline__ void go(double (&acc1)[R], const double (&T0)[R][L0P],
                                              uint32_t cp, uint32_t t1a, int cnt,
                                              uint32_t t1_row_bytes, uint32_t t1_r_bytes)
    {
#pragma unroll 1
        for (int j = 0; j < cnt; ++j) {
            double2 cf[K];
#pragma unroll
            for (int i = 0; i < K; ++i) cf[i] = lds128(cp + 16 * i);
            double t1v[R];
#pragma unroll
            for (int r = 0; r < R; ++r) t1v[r] = lds64(t1a + r * t1_r_bytes);
            double a[R][A];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if constexpr (A == 2) {
                    a[r][1] = cf[K - 1].y * T0[r][2 * K - 1];
                    a[r][0] = cf[K - 1].x * T0[r][2 * K - 2];
                } else {
                    a[r][0] = fma(cf[K - 1].x, T0[r][2 * K - 2], cf[K - 1].y * T0[r][2 * K - 1]);
                }
            }
#pragma unroll
            for (int i = K - 2; i >= 0; --i) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    a[r][A - 1] = fma(cf[i].y, T0[r][2 * i + 1], a[r][A - 1]);
                    a[r][0] = fma(cf[i].x, T0[r][2 * i], a[r][0]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if constexpr (A == 2) acc1[r] = fma(t1v[r], a[r][0] + a[r][1], acc1[r]);
                else acc1[r] = fma(t1v[r], a[r][0], acc1[r]);
            }
            cp += 16 * K;
            t1a += t1_row_bytes;
        }
    }
};

#define MPX_RUN_CASE(K)                                                                        \
    case (K):                                                                                  \
        if constexpr ((K) <= L0P / 2)                                                          \
            Run<L0P, R, A, ((K) <= L0P / 2 ? (K) : 1)>::go(acc1, T0, cp, t1a, cnt, t1_row_bytes, \
                                                           t1_r_bytes);                        \
        break;

template <int L0P, int R, int A>
__device__ __forceinline__ void run_dispatch(double (&acc1)[R], const double (&T0)[R][L0P],
                                             uint32_t cp, uint32_t t1a, int np, int cnt,
                                             uint32_t t1_row_bytes, uint32_t t1_r_bytes)
{
    switch (np) {
        MPX_RUN_CASE(1) MPX_RUN_CASE(2) MPX_RUN_CASE(3) MPX_RUN_CASE(4) MPX_RUN_CASE(5)
        MPX_RUN_CASE(6) MPX_RUN_CASE(7) MPX_RUN_CASE(8) MPX_RUN_CASE(9) MPX_RUN_CASE(10)
        MPX_RUN_CASE(11) MPX_RUN_CASE(12) MPX_RUN_CASE(13) MPX_RUN_CASE(14) MPX_RUN_CASE(15)
        MPX_RUN_CASE(16) MPX_RUN_CASE(17) MPX_RUN_CASE(18) MPX_RUN_CASE(19) MPX_RUN_CASE(20)
        MPX_RUN_CASE(21) MPX_RUN_CASE(22) MPX_RUN_CASE(23) MPX_RUN_CASE(24) MPX_RUN_CASE(25)
        MPX_RUN_CASE(26) MPX_RUN_CASE(27) MPX_RUN_CASE(28) MPX_RUN_CASE(29) MPX_RUN_CASE(30)
        MPX_RUN_CASE(31) MPX_RUN_CASE(32) MPX_RUN_CASE(33)
    default:
        break;
    }
}

// shared-memory carve-up (dynamic):
//   [0, 128)                     mbarriers: full[], empty[]
//   [128, 128 + S*STAGE_BYTES)   ring: per stage desc words then coefficients
//   then gp (m*gp_rows doubles), T1 ((max_exp1+1)*R*NT doubles),
//   upper-level state product/sum/coordinate (3 * U * R * NT doubles)
template <int L0P, int R, int A, int MINB>
__global__ void __launch_bounds__(256, MINB) eval_tree_kernel(const EvalParams p)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x;
    const int NT = blockDim.x;
    const int m = p.m;
    const int U = m > 2 ? m - 2 : 0;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    const int S = p.stages;
    uint64_t *empty = full + MPX_STAGES;
    unsigned char *ring = smem + 128;
    double *gps = reinterpret_cast<double *>(ring + S * STAGE_BYTES);
    double *T1s = gps + ((m * p.gp_rows + 1) & ~1);
    const int me1 = (m >= 2) ? p.max_exp[1] : 0;
    double *ups = T1s + (size_t)(me1 + 1) * R * NT;   // [3][U][R][NT]: product, sum, coordinate
    const size_t up_stride = (size_t)U * R * NT;

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
    // Producer = thread 0, never blocking: it tops the ring up whenever it polls a stage free.
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

    // ---- per-point tables -------------------------------------------------------------
    const int64_t tile0 = (int64_t)blockIdx.x * (R * NT);
    double T0[R][L0P];
    double acc1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int64_t pt = tile0 + (int64_t)r * NT + tid;
        if (pt >= p.P) pt = p.P - 1;               // tail threads recompute the last point
        const double *xp = p.xx + pt * p.ld_xx;
        const double x0 = xp[0];
        const int me0 = p.max_exp[0];
        T0[r][0] = 1.0;
#pragma unroll
        for (int j = 0; j < L0P - 1; ++j)
            T0[r][j + 1] = (j < me0) ? T0[r][j] * (x0 - gps[j]) : 0.0;
        if (m >= 2) {
            const double x1 = xp[1];
            double prod = 1.0;
            T1s[(size_t)r * NT + tid] = 1.0;
#pragma unroll 1
            for (int j = 0; j < me1; ++j) {
                prod *= x1 - gps[p.gp_rows + j];
                T1s[((size_t)(j + 1) * R + r) * NT + tid] = prod;
            }
        } else {
            T1s[(size_t)r * NT + tid] = 1.0;
        }
#pragma unroll 1
        for (int u = 0; u < U; ++u) {
            const double xu = xp[2 + u];
            double t = 1.0;
#pragma unroll 1
            for (int j = 0; j < p.start_exp[2 + u]; ++j) t *= xu - gps[(2 + u) * p.gp_rows + j];
            const size_t o = ((size_t)u * R + r) * NT + tid;
            ups[o] = t;
            ups[up_stride + o] = 0.0;
            ups[2 * up_stride + o] = xu;
        }
        acc1[r] = 0.0;
    }
    // (each thread only ever touches its own T1s / ups columns: no barrier needed)
    const uint32_t t1_base = smem_u32(T1s + tid);
    const uint32_t t1_r_bytes = (uint32_t)NT * 8u;            // T1 is [e][r][thread]
    const uint32_t t1_row_bytes = (uint32_t)R * t1_r_bytes;

    // ---- stream the program -------------------------------------------------------------
    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % S;
        const uint32_t ph = (uint32_t)(c / S) & 1u;
        while (next_load <= c) {
            // thread 0, rare: chunk c itself was never requested because the stage it needs
            // was still in use at every poll -- now it has to wait for that stage
            const int sp = next_load % S;
            mbar_wait(&empty[sp], (uint32_t)(next_load / S - 1) & 1u);
            issue_load(next_load);
            ++next_load;
        }
        mbar_wait(&full[s], ph);
        const uint32_t *dw = reinterpret_cast<const uint32_t *>(ring + s * STAGE_BYTES);
        uint32_t cp = smem_u32(ring + s * STAGE_BYTES + MPX_DESC_WORDS * 4);
        const int iend = MPX_DESC_HDR + (int)dw[0];
        uint32_t w = dw[MPX_DESC_HDR];
        for (int i = MPX_DESC_HDR; i < iend; ++i) {
            if ((i & 7) == 0 && next_load < p.n_chunks) {
                // thread 0 only: refill the oldest stage if every thread has released it
                const int sp = next_load % S;
                const uint32_t php = (uint32_t)(next_load / S - 1) & 1u;
                if (mbar_test(&empty[sp], php)) {
                    issue_load(next_load);
                    ++next_load;
                }
            }
            uint32_t w_next = dw[i + 1];     // next record, fetched a whole run ahead
            const int np = w & 63;
            const int e1 = (w >> 6) & 255;
            const int nf = (w >> 14) & 15;
            const int cnt = (int)(w >> 28) + 1;
            run_dispatch<L0P, R, A>(acc1, T0, cp, t1_base + (uint32_t)e1 * t1_row_bytes, np, cnt,
                                    t1_row_bytes, t1_r_bytes);
            cp += 16 * np * cnt;
            if (nf) {
                // fold finished subtrees upwards: levels 0..nf-1 (dimensions 2..nf+1)
                int adv = 1, cur = (w >> 19) & 255;
                const bool irregular = (w & MPX_D_IRREG) != 0;
                const bool last = (w & MPX_D_LAST) != 0;
                if (irregular) {
                    const uint32_t x = dw[++i];
                    cur = x & 255;
                    adv = (int)(x >> 8);
                }
                double low[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    low[r] = acc1[r];
                    acc1[r] = 0.0;
                }
#pragma unroll 1
                for (int u = 0; u < nf; ++u) {
                    double *ut = ups + (size_t)u * R * NT + tid;
                    if (u < nf - 1) {
                        // level complete as well: hand its sum upwards and restart it
                        const int st = irregular ? (int)dw[i + 1 + u] : 0;
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            low[r] = fma(ut[r * NT], low[r], ut[up_stride + r * NT]);
                            ut[up_stride + r * NT] = 0.0;
                            double tn = 1.0;
                            if (st) {
                                const double xu = ut[2 * up_stride + r * NT];
#pragma unroll 1
                                for (int j = 0; j < st; ++j) tn *= xu - gps[(2 + u) * p.gp_rows + j];
                            }
                            ut[r * NT] = tn;
                        }
                    } else {
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            double t = ut[r * NT];
                            ut[up_stride + r * NT] = fma(t, low[r], ut[up_stride + r * NT]);
                            if (!last) {
                                const double xu = ut[2 * up_stride + r * NT];
                                t *= xu - gps[(2 + u) * p.gp_rows + cur];
#pragma unroll 1
                                for (int j = 1; j < adv; ++j)
                                    t *= xu - gps[(2 + u) * p.gp_rows + cur + j];
                                ut[r * NT] = t;
                            }
                        }
                    }
                }
                if (irregular) {
                    i += nf - 1;
                    w_next = dw[i + 1];
                }
            }
            w = w_next;
        }
        mbar_arrive(&empty[s]);
    }

    // ---- result ---------------------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t pt = tile0 + (int64_t)r * NT + tid;
        if (pt < p.P) {
            const double res =
                (U > 0) ? ups[up_stride + ((size_t)(U - 1) * R + r) * NT + tid] : acc1[r];
            p.out[pt * p.ld_out + q] = res;
        }
    }
}

struct Variant {
    int l0p, r, nt;
    const void *fn;
};

template <int L0P, int R, int A, int MINB>
const void *kfn()
{
    return reinterpret_cast<const void *>(&eval_tree_kernel<L0P, R, A, MINB>);
}

size_t eval_smem_bytes(const mpx_plan *pl, int R, int NT, int stages)
{
    const int m = pl->m;
    const int U = m > 2 ? m - 2 : 0;
    const int me1 = m >= 2 ? pl->max_exp[1] : 0;
    size_t doubles = (size_t)((m * pl->gp_rows + 1) & ~1) + (size_t)(me1 + 1) * R * NT +
                     (size_t)3 * U * R * NT;
    return 128 + (size_t)stages * STAGE_BYTES + doubles * 8;
}

}  // namespace

int mpx_launch_eval(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx,
                    const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                    cudaStream_t stream)
{
    if (P == 0) return MPX_OK;
    const int L0 = pl->max_exp[0] + 1;
    // kernel variants: (dimension-0 table size, points per thread, min CTAs/SM)
    static const Variant variants[] = {
        {10, 4, 256, kfn<10, 4, 1, 1>()},
        {18, 4, 256, kfn<18, 4, 1, 1>()},
        {34, 2, 256, kfn<34, 2, 2, 1>()},
        {66, 1, 256, kfn<66, 1, 2, 1>()},
        // experimental alternatives, selected with MPX_EVAL_VARIANT=<index>
        {18, 4, 256, kfn<18, 4, 2, 1>()},   // 4: two accumulators per point
        {18, 2, 256, kfn<18, 2, 2, 2>()},   // 5: 2 points x 2 accumulators, 2 CTAs/SM
        {18, 3, 256, kfn<18, 3, 1, 1>()},   // 6
    };
    static const int env_variant = getenv("MPX_EVAL_VARIANT") ? atoi(getenv("MPX_EVAL_VARIANT")) : -1;
    const Variant *v = nullptr;
    for (const Variant &cand : variants)
        if (L0 <= cand.l0p) { v = &cand; break; }
    if (env_variant >= 0 && env_variant < (int)(sizeof(variants) / sizeof(variants[0])) &&
        L0 <= variants[env_variant].l0p)
        v = &variants[env_variant];
    if (!v) return mpx_fail(MPX_ERR_UNSUPPORTED, "dimension-0 exponent too large for the evaluation kernel");

    int smem_max = 0;
    MPX_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, pl->device));
    // tuning knobs for experiments (defaults are what the library ships with)
    static const int env_nt = getenv("MPX_EVAL_NT") ? atoi(getenv("MPX_EVAL_NT")) : 0;
    static const int env_stages = getenv("MPX_EVAL_STAGES") ? atoi(getenv("MPX_EVAL_STAGES")) : 0;
    int stages = env_stages > 0 ? env_stages : 4;
    if (stages > MPX_STAGES) stages = MPX_STAGES;
    if (stages < 2) stages = 2;
    int nt = env_nt > 0 ? env_nt : v->nt;
    while (nt > 32 && eval_smem_bytes(pl, v->r, nt, stages) > (size_t)smem_max) nt >>= 1;
    const size_t smem = eval_smem_bytes(pl, v->r, nt, stages);
    if (smem > (size_t)smem_max)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "tables of this index set do not fit shared memory");

    // pack the coefficients into the program image (plan scratch)
    const int64_t n_slots = (int64_t)pl->prog.n_chunks * MPX_COEF_CAP;
    int rc = mpx_ensure_scratch(pl, (size_t)n_slots * k * sizeof(double));
    if (rc != MPX_OK) return rc;
    {
        const int64_t total = n_slots * k;
        const int tb = 256;
        pack_coeffs_kernel<<<(unsigned)((total + tb - 1) / tb), tb, 0, stream>>>(
            coeffs, ld_c, k, pl->d_slot_src, n_slots, pl->d_scratch);
        MPX_CUDA(cudaGetLastError());
    }

    EvalParams prm;
    prm.xx = xx; prm.ld_xx = ld_xx; prm.P = P; prm.out = out; prm.ld_out = ld_out;
    prm.desc = pl->d_desc; prm.coef = pl->d_scratch; prm.gp = pl->d_gp;
    prm.n_chunks = pl->prog.n_chunks; prm.stages = stages; prm.m = pl->m; prm.gp_rows = pl->gp_rows;
    for (int i = 0; i < MPX_MAX_M; ++i) { prm.max_exp[i] = pl->max_exp[i]; prm.start_exp[i] = pl->start_exp[i]; }

    MPX_CUDA(cudaFuncSetAttribute(v->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t per_cta = (int64_t)v->r * nt;
    const int64_t tiles = (P + per_cta - 1) / per_cta;
    if (tiles > 0x7fffffffLL || k > 65535)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points or polynomials for one launch");
    dim3 grid((unsigned)tiles, (unsigned)k, 1);
    void *args[] = {(void *)&prm};
    MPX_CUDA(cudaLaunchKernel(v->fn, grid, dim3(nt, 1, 1), args, smem, stream));
    return MPX_OK;
}
