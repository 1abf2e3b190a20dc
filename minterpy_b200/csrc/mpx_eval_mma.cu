// K2: Newton evaluation of MANY polynomials at once on the FP64 tensor cores.
//   replaces the `newton_monomials @ coefficients` contraction of eval_newton_polynomials,
//   src/minterpy/utils.py:325-332 (monomial stage jit_compiled_utils.py:62-211), for k >= 8.
//
// out (P x k) = M (P x N) * C (N x k), where a row of M holds the N Newton monomials of one
// point.  M is never stored: every warp forms its A fragments -- monomial values straight in
// mma.sync m8n8k4 layout -- from per-point product tables T_i[e] in shared memory (the
// reference's `products_placeholder`, multiplied in the reference's order i = 0..m-1), and
// contracts them against coefficient tiles that all warps of the CTA stream through a
// cp.async.bulk + mbarrier ring.  sm_100a has no tcgen05 kind for f64; the FP64 tensor path is
// DMMA via mma.sync, measured at 37.1 TFLOP/s on the B200 (tools/fp64_peak.cu).
//
// Tiling: CTA = NW warps; warp tile = 32 points x 8*NT8 polynomials (4 x NT8 accumulator tiles
// of 8x8, 8*NT8 doubles per lane); K-chunk = 32 monomials (8 mma k-steps).  Lane l of a k-step
// owns monomial 4*ks + (l & 3): it loads that monomial's exponents once (8 bytes) and forms its
// value for the 4 points (l >> 2) + 8*mt of the warp tile, so monomial formation costs
// (m-1) DMUL per 8*NT8 tensor FMAs.
#include <cstdlib>

#include "mpx_common.h"
#include "mpx_ptx.h"

namespace {

constexpr int KC = 32;             // monomials per chunk
constexpr int MMA_STAGES = 3;

struct MmaParams {
    const double *xx;
    int64_t ld_xx;
    int64_t P;
    double *out;
    int64_t ld_out;
    int k;
    const uint64_t *exp64;   // n_chunks * KC packed exponents (8 bytes per monomial)
    const double *coef;      // [poly tile][chunk][KC][RS]
    const double *gp;        // m x gp_rows
    int n_chunks;
    int m;
    int gp_rows;
    int e1;                  // table entries per dimension (largest exponent + 1)
    int max_exp[8];
    int order[8];            // derivative order per dimension (all 0: plain evaluation)
};

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

// coefficient image: [tile][chunk][KC][RS], RS = 8*NT8 + 4 (bank-conflict-free fragment loads)
__global__ void pack_mma_kernel(const double *__restrict__ coeffs, int64_t ld_c, int k, int64_t N,
                                int n_chunks, int nt8, double *__restrict__ img)
{
    const int rs = 8 * nt8 + 4;
    const int64_t per_tile = (int64_t)n_chunks * KC * rs;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int tiles = (k + 8 * nt8 - 1) / (8 * nt8);
    if (t >= per_tile * tiles) return;
    const int tile = (int)(t / per_tile);
    const int64_t r = t % per_tile;
    const int64_t j = r / rs;          // monomial (row of the caller's matrix)
    const int c = (int)(r % rs);
    const int q = tile * 8 * nt8 + c;
    img[t] = (c < 8 * nt8 && q < k && j < N) ? coeffs[j * ld_c + q] : 0.0;
}

// Turns one column of product tables T[q] = prod_{t<q} (x - g_t) into its `order`-th derivative
// with respect to x, in place (polynomials/utils.py:8-112 computes the same numbers as
// order! * sum over all ways to drop `order` factors, itertools.combinations, at exponential
// cost).  With F_j[q] = (d/dx)^j T[q] / j!, the product rule on T[q+1] = (x - g_q) T[q] gives
//     F_j[q+1] = (x - g_q) F_j[q] + F_{j-1}[q],     F_j[0] = 0 for j >= 1,
// so pass j rewrites the column from F_{j-1} to F_j walking q upwards (the old entry of row
// q+1 is read just before it is overwritten).  Entries below `order` come out as exact zeros
// and entry `order` as exactly order!, like the reference's table.
template <int PS>
__device__ __forceinline__ void derivative_table(double *T, double x, const double *g, int max_exp,
                                                 int order)
{
    if (order > max_exp) {
        for (int q = 0; q <= max_exp; ++q) T[(size_t)q * PS] = 0.0;
        return;
    }
    for (int pass = 1; pass <= order; ++pass) {
        double fj = 0.0, old = T[0];
        T[0] = 0.0;
        for (int q = 0; q < max_exp; ++q) {
            const double old_next = T[(size_t)(q + 1) * PS];
            fj = __fma_rn(x - g[q], fj, old);
            T[(size_t)(q + 1) * PS] = fj;
            old = old_next;
        }
    }
    double fact = 1.0;
    for (int t = 2; t <= order; ++t) fact *= (double)t;
    for (int q = order; q <= max_exp; ++q) T[(size_t)q * PS] *= fact;
}

// MP = number of dimensions the monomial product is unrolled for (>= m; dimensions m..MP-1
// read a table row that holds 1.0, so the k-step stays one branch-free basic block and the
// compiler can interleave the table reads of one k-step with the tensor ops of another)
template <int NT8, int NW, int MP>
__global__ void __launch_bounds__(32 * NW, 1) eval_mma_kernel(const MmaParams p)
{
    constexpr int RS = 8 * NT8 + 4;
    constexpr int STAGE_BYTES = KC * 8 + KC * RS * 8;
    constexpr int NTH = 32 * NW;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + MMA_STAGES;
    unsigned char *ring = smem + 128;
    double *gps = reinterpret_cast<double *>(ring + MMA_STAGES * STAGE_BYTES);
    double *tab = gps + ((p.m * p.gp_rows + 1) & ~1);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = p.m, E1 = p.e1;
    const int tile = blockIdx.y;
    const double *coef_img = p.coef + (size_t)tile * p.n_chunks * (KC * RS);

    if (tid == 0) {
        for (int s = 0; s < MMA_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NTH);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < m * p.gp_rows; i += NTH) gps[i] = p.gp[i];
    __syncthreads();
    int next_load = 0;
    auto issue_load = [&](int cn) {
        const int sp = cn % MMA_STAGES;
        unsigned char *dst = ring + sp * STAGE_BYTES;
        mbar_arrive_expect_tx(&full[sp], STAGE_BYTES);
        bulk_g2s(dst, p.exp64 + (size_t)cn * KC, KC * 8, &full[sp]);
        bulk_g2s(dst + KC * 8, coef_img + (size_t)cn * (KC * RS), KC * RS * 8, &full[sp]);
    };
    if (tid == 0) {
        const int pre = p.n_chunks < MMA_STAGES ? p.n_chunks : MMA_STAGES;
        for (; next_load < pre; ++next_load) issue_load(next_load);
    } else {
        next_load = p.n_chunks;
    }

    // product tables of this CTA's points: tab[dim][e][point], thread = point.  The point
    // stride PS = points + 4 keeps the fragment reads below conflict free: the 4 monomials of
    // a k-step usually differ in one exponent only (rows 8 words apart mod 32), the 4 points of
    // a half-warp are 2 words apart.
    constexpr int PS = NTH + 4;
    const int64_t p0 = (int64_t)blockIdx.x * NTH;
    {
        int64_t pt = p0 + tid;
        if (pt >= p.P) pt = p.P - 1;
        const double *xp = p.xx + pt * p.ld_xx;
        for (int i = 0; i < MP; ++i) {
            double *T = tab + (size_t)i * E1 * PS + tid;
            T[0] = 1.0;
            if (i >= m) continue;                        // padding dimension: exponent is always 0
            const double x = xp[i];
            double prod = 1.0;
            for (int j = 0; j < p.max_exp[i]; ++j) {     // jit_compiled_utils.py:102-112
                prod *= x - gps[i * p.gp_rows + j];
                T[(size_t)(j + 1) * PS] = prod;
            }
            for (int j = p.max_exp[i] + 1; j < E1; ++j) T[(size_t)j * PS] = 0.0;
            if (p.order[i] > 0) derivative_table<PS>(T, x, gps + i * p.gp_rows, p.max_exp[i], p.order[i]);
        }
    }
    __syncthreads();

    double acc[4][NT8][2];
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT8; ++nt) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;

    const int jl = lane & 3, g = lane >> 2;
    const double *tw = tab + warp * 32 + g;                  // point g of m-tile 0; +8 per m-tile

    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % MMA_STAGES;
        const uint32_t ph = (uint32_t)(c / MMA_STAGES) & 1u;
        if (next_load < p.n_chunks) {
            for (;;) {
                const int sp = next_load % MMA_STAGES;
                const uint32_t php = (uint32_t)(next_load / MMA_STAGES - 1) & 1u;
                if (next_load <= c) mbar_wait(&empty[sp], php);
                else if (!mbar_test(&empty[sp], php)) break;
                issue_load(next_load);
                if (++next_load >= p.n_chunks) break;
            }
        }
        mbar_wait(&full[s], ph);
        const uint64_t *ex = reinterpret_cast<const uint64_t *>(ring + s * STAGE_BYTES);
        const double *cf = reinterpret_cast<const double *>(ring + s * STAGE_BYTES + KC * 8);
#pragma unroll 2
        for (int ks = 0; ks < KC / 4; ++ks) {
            const int j = 4 * ks + jl;
            const uint2 e = *reinterpret_cast<const uint2 *>(ex + j);   // 8 exponents, one byte each
            // monomial value for the 4 points of this lane, reference order i = 0..m-1
            double a[4];
            {
                const double *row = tw + (e.x & 255u) * (uint32_t)PS;
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) a[mt] = row[8 * mt];
            }
#pragma unroll
            for (int i = 1; i < MP; ++i) {
                const uint32_t ei = ((i < 4 ? e.x : e.y) >> (8 * (i & 3))) & 255u;
                const double *row = tw + ((uint32_t)(i * E1) + ei) * (uint32_t)PS;
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) a[mt] *= row[8 * mt];
            }
            const double *brow = cf + j * RS + g;
#pragma unroll
            for (int nt = 0; nt < NT8; ++nt) {
                const double b = brow[8 * nt];
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], b);
            }
        }
        mbar_arrive(&empty[s]);
    }

    // C fragment: row = lane >> 2, columns 2*(lane & 3) + {0,1}
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
        const int64_t pt = p0 + warp * 32 + mt * 8 + g;
        if (pt >= p.P) continue;
#pragma unroll
        for (int nt = 0; nt < NT8; ++nt) {
            const int q = tile * 8 * NT8 + nt * 8 + 2 * jl;
            double *o = p.out + pt * p.ld_out + q;
            if (q < p.k) o[0] = acc[mt][nt][0];
            if (q + 1 < p.k) o[1] = acc[mt][nt][1];
        }
    }
}

template <int NT8, int NW, int MP>
int launch_variant(const MmaParams &prm, int tiles, size_t smem, cudaStream_t stream)
{
    auto fn = &eval_mma_kernel<NT8, NW, MP>;
    MPX_CUDA(mpx_allow_max_smem(fn));
    const int64_t ptiles = (prm.P + 32 * NW - 1) / (32 * NW);
    if (ptiles > 0x7fffffffLL || tiles > 65535)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points or polynomials for one launch");
    fn<<<dim3((unsigned)ptiles, (unsigned)tiles, 1), 32 * NW, smem, stream>>>(prm);
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}

template <int NT8, int NW>
int launch_mp(int mp, const MmaParams &prm, int tiles, size_t smem, cudaStream_t stream)
{
    switch (mp) {
    case 2: return launch_variant<NT8, NW, 2>(prm, tiles, smem, stream);
    case 3: return launch_variant<NT8, NW, 3>(prm, tiles, smem, stream);
    case 4: return launch_variant<NT8, NW, 4>(prm, tiles, smem, stream);
    case 5: return launch_variant<NT8, NW, 5>(prm, tiles, smem, stream);
    case 6: return launch_variant<NT8, NW, 6>(prm, tiles, smem, stream);
    default: return launch_variant<NT8, NW, 8>(prm, tiles, smem, stream);
    }
}

// polynomials per CTA tile for k polynomials: 8, 16, 32 or 64
int mma_nt8_for(int k) { return k <= 8 ? 1 : k <= 16 ? 2 : k <= 32 ? 4 : 8; }
constexpr size_t MMA_SMEM_LIMIT = 227 * 1024;
int mma_padded_dims(int m) { return m <= 2 ? 2 : m <= 6 ? m : 8; }
size_t mma_smem_bytes(const mpx_plan *pl, int nt8, int nw)
{
    int e1 = 1;
    for (int i = 0; i < pl->m; ++i) e1 = std::max(e1, pl->max_exp[i] + 1);
    const int rs = 8 * nt8 + 4;
    return (size_t)128 + (size_t)MMA_STAGES * (KC * 8 + KC * rs * 8) +
           (size_t)((pl->m * pl->gp_rows + 1) & ~1) * 8 +
           (size_t)mma_padded_dims(pl->m) * e1 * (32 * nw + 4) * 8;
}

}  // namespace


// Smallest k that goes to this kernel when a tree kernel could take the plan as well.
// Crossover against the tree kernel run once per polynomial (measured in round 2 with the
// uniform-control tree kernel, 1e6 points, tools/bench_k_sweep.py; this kernel costs the same for
// every k of a tile size 8 / 16 / 32): cfg2 (10 monomials per leaf) 2.33 ms per polynomial against
// 31.8 / 42.6 / 61.6 ms -- the tree kernel wins up to k = 26; cfg5 (33 per leaf, tensor-core tree
// kernel) 2.98 ms against 41.8 / 54.9 / 114.9 -- up to k = 32 and beyond; cfg4 (7 per leaf) 7.3 ms
// against 84 / 109 / 159 -- up to k = 12 (and 13, 14, 20 by a few per cent: ignored); cfg3 (2.7
// per leaf) 3.3 ms against 13.4 / 16.4 / 21.1 -- level at k = 4, this kernel from there.
int mpx_mma_min_k(const mpx_plan *pl)
{
#ifdef MPX_TUNING
    static const int env_min_k = getenv("MPX_EVAL_MMA_MIN_K") ? atoi(getenv("MPX_EVAL_MMA_MIN_K")) : 0;
    if (env_min_k > 0) return env_min_k;
#endif
    if (pl->tree_ok && pl->prog.n_leaves > 0) {
        if (pl->N >= 16 * pl->prog.n_leaves) return 33;
        if (pl->N >= 8 * pl->prog.n_leaves) return 27;
        if (pl->N >= 6 * pl->prog.n_leaves) return 13;
        if (pl->N < 4 * pl->prog.n_leaves) return 4;
    }
    return 8;
}

bool mpx_mma_eligible(const mpx_plan *pl, int k)
{
    // the product tables of at least 64 points must fit shared memory, else the tree kernel
    // runs; plans the tree kernel cannot take (dimension-0 exponent > 64) come here for every k
    return (k >= mpx_mma_min_k(pl) || !pl->tree_ok) && pl->m <= 8 && pl->d_exp64 != nullptr &&
           mma_smem_bytes(pl, mma_nt8_for(k), 2) <= MMA_SMEM_LIMIT;
}

size_t mpx_mma_scratch_bytes(const mpx_plan *pl, int k)
{
    const int nt8 = mma_nt8_for(k);
    const int tiles = (k + 8 * nt8 - 1) / (8 * nt8);
    return (size_t)tiles * pl->mma_chunks * KC * (8 * nt8 + 4) * sizeof(double);
}

bool mpx_mma_fits(const mpx_plan *pl, int k)
{
    return pl->m <= 8 && pl->d_exp64 != nullptr &&
           mma_smem_bytes(pl, mma_nt8_for(k), 2) <= MMA_SMEM_LIMIT;
}

int mpx_launch_eval_mma(mpx_plan *pl, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                        const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                        cudaStream_t stream, const int *orders_h)
{
    const int nt8 = mma_nt8_for(k);
    const int tiles = (k + 8 * nt8 - 1) / (8 * nt8);
    const int rs = 8 * nt8 + 4;
    if (coeffs) {
        int rc = mpx_image_acquire(img, mpx_mma_scratch_bytes(pl, k), stream);
        if (rc != MPX_OK) return rc;
        const int64_t total = (int64_t)tiles * pl->mma_chunks * KC * rs;
        pack_mma_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
            coeffs, ld_c, k, pl->N, pl->mma_chunks, nt8, img->ptr);
        MPX_CUDA(cudaGetLastError());
    }
    if (P == 0) return coeffs ? mpx_image_release(img, stream) : MPX_OK;

    MmaParams prm;
    prm.xx = xx; prm.ld_xx = ld_xx; prm.P = P; prm.out = out; prm.ld_out = ld_out; prm.k = k;
    prm.exp64 = pl->d_exp64; prm.coef = img->ptr; prm.gp = pl->d_gp;
    prm.n_chunks = pl->mma_chunks; prm.m = pl->m; prm.gp_rows = pl->gp_rows;
    int e1 = 1;
    for (int i = 0; i < pl->m; ++i) { prm.max_exp[i] = pl->max_exp[i]; e1 = std::max(e1, pl->max_exp[i] + 1); }
    for (int i = pl->m; i < 8; ++i) prm.max_exp[i] = 0;
    for (int i = 0; i < 8; ++i) prm.order[i] = (orders_h && i < pl->m) ? orders_h[i] : 0;
    prm.e1 = e1;

    const int nw = mma_smem_bytes(pl, nt8, 8) <= MMA_SMEM_LIMIT ? 8 : 2;
    const size_t smem = mma_smem_bytes(pl, nt8, nw);
    const int mp = mma_padded_dims(pl->m);
#define MPX_MMA_CASE(NT8, NW)                                              \
    if (nt8 == NT8 && nw == NW) {                                          \
        int rc = launch_mp<NT8, NW>(mp, prm, tiles, smem, stream);         \
        return rc != MPX_OK ? rc : mpx_image_release(img, stream);         \
    }
    MPX_MMA_CASE(8, 8) MPX_MMA_CASE(8, 2) MPX_MMA_CASE(4, 8) MPX_MMA_CASE(4, 2)
    MPX_MMA_CASE(2, 8) MPX_MMA_CASE(2, 2) MPX_MMA_CASE(1, 8) MPX_MMA_CASE(1, 2)
#undef MPX_MMA_CASE
    return mpx_fail(MPX_ERR_UNSUPPORTED, "no tensor-core variant for this shape");
}
