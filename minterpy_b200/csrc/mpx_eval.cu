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

template <int L0P, int R, int I>
struct LeafPairs {
    // pairs I-1 .. 0 are handled by the switch in leaf_sum; this helper does pair I
    static __device__ __forceinline__ void run(double (&a)[R], const double (&T0)[R][L0P],
                                               const double2 *cp)
    {
        const double2 c = cp[I];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            a[r] = fma(c.y, T0[r][2 * I + 1], a[r]);
            a[r] = fma(c.x, T0[r][2 * I], a[r]);
        }
    }
};

#define MPX_LEAF_CASE(I)                                                          \
    case (I) + 1:                                                                 \
        if constexpr ((I) < L0P / 2) LeafPairs<L0P, R, ((I) < L0P / 2 ? (I) : 0)>::run(a, T0, cp); \
        /* fallthrough */

template <int L0P, int R>
__device__ __forceinline__ void leaf_sum(double (&a)[R], const double (&T0)[R][L0P],
                                         const double2 *cp, int np)
{
    switch (np) {
        MPX_LEAF_CASE(32) MPX_LEAF_CASE(31) MPX_LEAF_CASE(30) MPX_LEAF_CASE(29)
        MPX_LEAF_CASE(28) MPX_LEAF_CASE(27) MPX_LEAF_CASE(26) MPX_LEAF_CASE(25)
        MPX_LEAF_CASE(24) MPX_LEAF_CASE(23) MPX_LEAF_CASE(22) MPX_LEAF_CASE(21)
        MPX_LEAF_CASE(20) MPX_LEAF_CASE(19) MPX_LEAF_CASE(18) MPX_LEAF_CASE(17)
        MPX_LEAF_CASE(16) MPX_LEAF_CASE(15) MPX_LEAF_CASE(14) MPX_LEAF_CASE(13)
        MPX_LEAF_CASE(12) MPX_LEAF_CASE(11) MPX_LEAF_CASE(10) MPX_LEAF_CASE(9)
        MPX_LEAF_CASE(8) MPX_LEAF_CASE(7) MPX_LEAF_CASE(6) MPX_LEAF_CASE(5)
        MPX_LEAF_CASE(4) MPX_LEAF_CASE(3) MPX_LEAF_CASE(2) MPX_LEAF_CASE(1)
        MPX_LEAF_CASE(0)
    default:
        break;
    }
}

// shared-memory carve-up (dynamic):
//   [0, 64)                      mbarriers: full[S], empty[S]
//   [64, 64 + S*STAGE_BYTES)     ring: per stage desc words then coefficients
//   then gp (m*gp_rows doubles), T1 ((max_exp1+1)*R*NT doubles),
//   upper-level state x/t/acc (3 * U * R * NT doubles)
template <int L0P, int R, int MINB>
__global__ void __launch_bounds__(256, MINB) eval_tree_kernel(const EvalParams p)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x;
    const int NT = blockDim.x;
    const int m = p.m;
    const int U = m > 2 ? m - 2 : 0;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + MPX_STAGES;
    unsigned char *ring = smem + 64;
    double *gps = reinterpret_cast<double *>(ring + MPX_STAGES * STAGE_BYTES);
    double *T1s = gps + ((m * p.gp_rows + 1) & ~1);
    const int me1 = (m >= 2) ? p.max_exp[1] : 0;
    double *ups = T1s + (size_t)(me1 + 1) * R * NT;   // [3][U][R][NT]: x, t, acc
    const size_t up_stride = (size_t)U * R * NT;

    const int q = blockIdx.y;                                   // polynomial
    const double *coef_img = p.coef + (size_t)q * p.n_chunks * MPX_COEF_CAP;

    if (tid == 0) {
        for (int s = 0; s < MPX_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NT);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < m * p.gp_rows; i += NT) gps[i] = p.gp[i];
    __syncthreads();
    if (tid == 0) {
        const int pre = p.n_chunks < MPX_STAGES ? p.n_chunks : MPX_STAGES;
        for (int c = 0; c < pre; ++c) {
            unsigned char *dst = ring + c * STAGE_BYTES;
            mbar_arrive_expect_tx(&full[c], STAGE_BYTES);
            bulk_g2s(dst, p.desc + (size_t)c * MPX_DESC_WORDS, MPX_DESC_WORDS * 4, &full[c]);
            bulk_g2s(dst + MPX_DESC_WORDS * 4, coef_img + (size_t)c * MPX_COEF_CAP,
                     MPX_COEF_CAP * 8, &full[c]);
        }
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
            ups[o] = xu;
            ups[up_stride + o] = t;
            ups[2 * up_stride + o] = 0.0;
        }
        acc1[r] = 0.0;
    }
    // (each thread only ever touches its own T1s / ups columns: no barrier needed)

    // ---- stream the program -------------------------------------------------------------
    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % MPX_STAGES;
        const uint32_t ph = (uint32_t)(c / MPX_STAGES) & 1u;
        if (tid == 0 && c >= 1 && c - 1 + MPX_STAGES < p.n_chunks) {
            // refill the stage chunk c-1 lived in, once every thread has released it
            const int sp = (c - 1) % MPX_STAGES;
            const uint32_t php = (uint32_t)((c - 1) / MPX_STAGES) & 1u;
            mbar_wait(&empty[sp], php);
            const int cn = c - 1 + MPX_STAGES;
            unsigned char *dst = ring + sp * STAGE_BYTES;
            mbar_arrive_expect_tx(&full[sp], STAGE_BYTES);
            bulk_g2s(dst, p.desc + (size_t)cn * MPX_DESC_WORDS, MPX_DESC_WORDS * 4, &full[sp]);
            bulk_g2s(dst + MPX_DESC_WORDS * 4, coef_img + (size_t)cn * MPX_COEF_CAP,
                     MPX_COEF_CAP * 8, &full[sp]);
        }
        mbar_wait(&full[s], ph);
        const uint32_t *dw = reinterpret_cast<const uint32_t *>(ring + s * STAGE_BYTES);
        const double2 *cp =
            reinterpret_cast<const double2 *>(ring + s * STAGE_BYTES + MPX_DESC_WORDS * 4);
        const int nw = (int)dw[0];
        for (int i = MPX_DESC_HDR; i < MPX_DESC_HDR + nw; ++i) {
            const uint32_t w = dw[i];
            const int np = w & 63;
            const int e1 = (w >> 6) & 255;
            const int nf = (w >> 14) & 15;
            double a[R];
#pragma unroll
            for (int r = 0; r < R; ++r) a[r] = 0.0;
            leaf_sum<L0P, R>(a, T0, cp, np);
            cp += np;
            const double *t1 = T1s + (size_t)e1 * R * NT + tid;
#pragma unroll
            for (int r = 0; r < R; ++r) acc1[r] = fma(t1[(size_t)r * NT], a[r], acc1[r]);
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
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    double low = acc1[r];
                    acc1[r] = 0.0;
#pragma unroll 1
                    for (int u = 0; u < nf; ++u) {
                        const size_t o = ((size_t)u * R + r) * NT + tid;
                        double t = ups[up_stride + o];
                        double av = fma(t, low, ups[2 * up_stride + o]);
                        if (u < nf - 1) {
                            // this level is complete as well: hand its sum upwards, restart it
                            low = av;
                            ups[2 * up_stride + o] = 0.0;
                            double tn = 1.0;
                            if (irregular) {
                                const int st = (int)dw[i + 1 + u];
                                const double xu = ups[o];
#pragma unroll 1
                                for (int j = 0; j < st; ++j) tn *= xu - gps[(2 + u) * p.gp_rows + j];
                            }
                            ups[up_stride + o] = tn;
                        } else {
                            ups[2 * up_stride + o] = av;
                            if (!last) {
                                const double xu = ups[o];
#pragma unroll 1
                                for (int j = 0; j < adv; ++j)
                                    t *= xu - gps[(2 + u) * p.gp_rows + cur + j];
                                ups[up_stride + o] = t;
                            }
                        }
                    }
                }
                if (irregular) i += nf - 1;
            }
        }
        mbar_arrive(&empty[s]);
    }

    // ---- result ---------------------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t pt = tile0 + (int64_t)r * NT + tid;
        if (pt < p.P) {
            const double res =
                (U > 0) ? ups[2 * up_stride + ((size_t)(U - 1) * R + r) * NT + tid] : acc1[r];
            p.out[pt * p.ld_out + q] = res;
        }
    }
}

struct Variant {
    int l0p, r, nt;
    const void *fn;
};

template <int L0P, int R, int MINB>
const void *kfn()
{
    return reinterpret_cast<const void *>(&eval_tree_kernel<L0P, R, MINB>);
}

size_t eval_smem_bytes(const mpx_plan *pl, int R, int NT)
{
    const int m = pl->m;
    const int U = m > 2 ? m - 2 : 0;
    const int me1 = m >= 2 ? pl->max_exp[1] : 0;
    size_t doubles = (size_t)((m * pl->gp_rows + 1) & ~1) + (size_t)(me1 + 1) * R * NT +
                     (size_t)3 * U * R * NT;
    return 64 + (size_t)MPX_STAGES * STAGE_BYTES + doubles * 8;
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
        {10, 2, 256, kfn<10, 2, 2>()},
        {18, 2, 256, kfn<18, 2, 2>()},
        {34, 2, 256, kfn<34, 2, 1>()},
        {66, 1, 256, kfn<66, 1, 1>()},
    };
    const Variant *v = nullptr;
    for (const Variant &cand : variants)
        if (L0 <= cand.l0p) { v = &cand; break; }
    if (!v) return mpx_fail(MPX_ERR_UNSUPPORTED, "dimension-0 exponent too large for the evaluation kernel");

    int smem_max = 0;
    MPX_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, pl->device));
    int nt = v->nt;
    while (nt > 32 && eval_smem_bytes(pl, v->r, nt) > (size_t)smem_max) nt >>= 1;
    const size_t smem = eval_smem_bytes(pl, v->r, nt);
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
    prm.n_chunks = pl->prog.n_chunks; prm.m = pl->m; prm.gp_rows = pl->gp_rows;
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
