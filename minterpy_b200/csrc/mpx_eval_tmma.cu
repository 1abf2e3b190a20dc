// K1T: Newton evaluation, tree form, with the leaf sums on the FP64 tensor cores.
//   replaces eval_newton_polynomials (src/minterpy/utils.py:247-343, inner loops
//   src/minterpy/jit_compiled_utils.py:62-126) for complete index sets and k < 8.
//
// p(x) = sum_{e_{m-1}} T_{m-1}[..] ( ... sum_{e_1} T_1[e_1] ( sum_{e_0} c_e T_0[e_0] ) ) as in
// mpx_eval.cu, but the innermost sums of 8 leaves at a time are one dense contraction
//   C[point][leaf] = sum_{e0} T_0[point][e0] * c[leaf][e0]
// = [points x e0] * [e0 x leaves], run as mma.sync m8n8k4 f64 (DMMA):
//   * A fragments = T_0 of the warp's 32 points, built once per CTA and kept in registers
//     (KS0 x 4 doubles per lane instead of the whole table per point);
//   * B fragments = coefficients, stored by the plan in fragment order (one coalesced,
//     conflict-free LDS.64 per k-step), streamed through a cp.async.bulk + mbarrier ring;
//   * epilogue per tile: acc1 += T_1[e1(column)] * C, columns of the tile may belong to the
//     tail of one (e2..) group and the head of the next (two partial sums per lane);
//   * when a group ends, the four lanes that share a point add their partial sums (shuffles)
//     and the sum is folded into dimension 2 (registers) and further up (shared memory).
// One DMMA issues 256 FMAs, so the kernel needs ~8x fewer instruction issues per FMA than the
// DFMA tree kernel and far fewer registers; the price is zero padding of the tiles (leaf
// lengths rounded up to 4 per tile): 1.36 tensor FMAs per monomial at BASELINE config 2.
#include <cstdlib>

#include "mpx_common.h"
#include "mpx_ptx.h"

namespace {

constexpr int T_STAGES = 3;
constexpr int T_STAGE_BYTES = MPX_T_DESC_WORDS * 4 + MPX_T_COEF_CAP * 8;

struct TmmaParams {
    const double *xx;
    int64_t ld_xx;
    int64_t P;
    double *out;
    int64_t ld_out;
    const uint32_t *desc;    // n_chunks * MPX_T_DESC_WORDS
    const double *coef;      // k * n_chunks * MPX_T_COEF_CAP (packed per call)
    const double *gp;        // m x gp_rows, dimension-major
    int n_chunks;
    int m;
    int gp_rows;
    int max_exp[8];
};

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

__global__ void pack_tile_coeffs_kernel(const double *__restrict__ coeffs, int64_t ld_c, int k,
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

__device__ __forceinline__ double quad_sum(double v)
{
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}

// shared memory (dynamic): [0,128) mbarriers | ring T_STAGES x T_STAGE_BYTES | gp (m x gp_rows)
// | T1 [(max_exp1+2)][PS] (last row zero) | x of dimensions 2.. [U][PS] | running product and partial sum of
// dimensions 3.. [2][U-1][PS];  PS = points + 2 (bank spread of the T1 rows)
// MT = 8-point m-tiles per warp (4: fewest issues per FMA; 2: half the A-fragment registers,
// more resident warps)
template <int KS0, int NW, int MT, int MINB>
__global__ void __launch_bounds__(32 * NW, MINB) eval_tmma_kernel(const TmmaParams p)
{
    constexpr int NTH = 32 * NW;
    constexpr int PTS = NW * 8 * MT;          // points per CTA
    // row stride of the point-indexed tables: PS*2 words = 4 (mod 32), so the rows e1 + 2*jl + h
    // read by the four lane groups of a half-warp fall on disjoint banks
    constexpr int PS = PTS + 2;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + T_STAGES;
    unsigned char *ring = smem + 128;
    double *gps = reinterpret_cast<double *>(ring + T_STAGES * T_STAGE_BYTES);
    const int m = p.m;
    const int U = m - 2;
    const int me1 = p.max_exp[1];
    double *T1s = gps + ((m * p.gp_rows + 1) & ~1);
    double *xs = T1s + (size_t)(me1 + 2) * PS;                 // [U][PS]; T1 row me1+1 is all zero
    double *ups = xs + (size_t)(U > 0 ? U : 0) * PS;           // [2][U-1][PS]
    const size_t up_stride = (size_t)(U > 1 ? U - 1 : 0) * PS;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int jl = lane & 3, g = lane >> 2;
    const int q = blockIdx.y;
    const double *coef_img = p.coef + (size_t)q * p.n_chunks * MPX_T_COEF_CAP;

    if (tid == 0) {
        for (int s = 0; s < T_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NTH);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < m * p.gp_rows; i += NTH) gps[i] = p.gp[i];
    __syncthreads();
    int next_load = 0;
    auto issue_load = [&](int cn) {
        const int sp = cn % T_STAGES;
        unsigned char *dst = ring + sp * T_STAGE_BYTES;
        mbar_arrive_expect_tx(&full[sp], T_STAGE_BYTES);
        bulk_g2s(dst, p.desc + (size_t)cn * MPX_T_DESC_WORDS, MPX_T_DESC_WORDS * 4, &full[sp]);
        bulk_g2s(dst + MPX_T_DESC_WORDS * 4, coef_img + (size_t)cn * MPX_T_COEF_CAP,
                 MPX_T_COEF_CAP * 8, &full[sp]);
    };
    if (tid == 0) {
        const int pre = p.n_chunks < T_STAGES ? p.n_chunks : T_STAGES;
        for (; next_load < pre; ++next_load) issue_load(next_load);
    } else {
        next_load = p.n_chunks;
    }

    // ---- per-point tables: thread = point for the shared-memory ones ----------------------
    const int64_t p0 = (int64_t)blockIdx.x * PTS;
    if (tid < PTS) {
        int64_t pt = p0 + tid;
        if (pt >= p.P) pt = p.P - 1;
        const double *xp = p.xx + pt * p.ld_xx;
        const double x1 = xp[1];
        double prod = 1.0;
        T1s[tid] = 1.0;
        for (int j = 0; j < me1; ++j) {              // jit_compiled_utils.py:102-112
            prod *= x1 - gps[p.gp_rows + j];
            T1s[(size_t)(j + 1) * PS + tid] = prod;
        }
        T1s[(size_t)(me1 + 1) * PS + tid] = 0.0;     // row for columns that do not take part
        for (int u = 0; u < U; ++u) {
            xs[(size_t)u * PS + tid] = xp[2 + u];
            if (u >= 1) {
                ups[(size_t)(u - 1) * PS + tid] = 1.0;
                ups[up_stride + (size_t)(u - 1) * PS + tid] = 0.0;
            }
        }
    }
    // A fragments: T0[point 8*mt + g of this warp][e0 = 4*s + jl]
    double a[KS0][MT];
    int64_t ptq[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        int64_t pt = p0 + warp * (8 * MT) + mt * 8 + g;
        ptq[mt] = pt;
        if (pt >= p.P) pt = p.P - 1;
        const double x0 = p.xx[pt * p.ld_xx];
        const int me0 = p.max_exp[0];
        double cur = 1.0;
#pragma unroll
        for (int j = 0; j < 4 * KS0; ++j) {
            if ((j & 3) == jl) a[j >> 2][mt] = cur;
            cur = (j < me0) ? cur * (x0 - gps[j]) : 0.0;
        }
    }
    __syncthreads();

    const int pbase = warp * (8 * MT) + g;                 // shared-memory column of m-tile 0; +8 per tile
    double acc_cur[MT], t2[MT], a2[MT];
    const uint32_t t1_lane = smem_u32(T1s + pbase);      // T1[0][point of m-tile 0]; +64 B per tile
    const uint32_t zrow = (uint32_t)(me1 + 1);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        acc_cur[mt] = 0.0;
        t2[mt] = 1.0;
        a2[mt] = 0.0;
    }

    // ---- stream the tile program ------------------------------------------------------------
    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % T_STAGES;
        const uint32_t ph = (uint32_t)(c / T_STAGES) & 1u;
        if (next_load < p.n_chunks) {
            // thread 0: refill released stages; wait only if chunk c itself is still missing
            for (;;) {
                const int sp = next_load % T_STAGES;
                const uint32_t php = (uint32_t)(next_load / T_STAGES - 1) & 1u;
                if (next_load <= c) mbar_wait(&empty[sp], php);
                else if (!mbar_test(&empty[sp], php)) break;
                issue_load(next_load);
                if (++next_load >= p.n_chunks) break;
            }
        }
        mbar_wait(&full[s], ph);
        const uint32_t *dw = reinterpret_cast<const uint32_t *>(ring + s * T_STAGE_BYTES);
        uint32_t cf = smem_u32(ring + s * T_STAGE_BYTES + MPX_T_DESC_WORDS * 4) + 16u * lane;
        const int iend = MPX_DESC_HDR + (int)dw[0];
        for (int i = MPX_DESC_HDR; i < iend; i += 2) {
            if (next_load < p.n_chunks) {            // thread 0, between tiles
                const int sp = next_load % T_STAGES;
                if (mbar_test(&empty[sp], (uint32_t)(next_load / T_STAGES - 1) & 1u)) {
                    issue_load(next_load);
                    ++next_load;
                }
            }
            const uint32_t w0 = dw[i], w1 = dw[i + 1];
            const int ks = w0 & 31;
            const int ncur = (w0 >> 5) & 15, nnext = (w0 >> 9) & 15;
            const int e1 = (w0 >> 13) & 255;
            double cacc[MT][2];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) cacc[mt][0] = cacc[mt][1] = 0.0;
#pragma unroll
            for (int sp = 0; sp < (KS0 + 1) / 2; ++sp) {
                if (2 * sp < ks) {
                    // B fragments of two k-steps in one 16-byte load (conflict free)
                    const double2 b = lds128(cf + 512 * sp);
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) dmma884(cacc[mt][0], cacc[mt][1], a[2 * sp][mt], b.x);
                    {
                        if (2 * sp + 1 < KS0 && 2 * sp + 1 < ks) {
#pragma unroll
                            for (int mt = 0; mt < MT; ++mt)
                                dmma884(cacc[mt][0], cacc[mt][1], a[(2 * sp + 1 < KS0) ? 2 * sp + 1 : 0][mt], b.y);
                        }
                    }
                }
            }
            cf += 512u * (uint32_t)((ks + 1) >> 1);
            // epilogue: this lane holds columns 2*jl and 2*jl+1 of its 4 points.  Columns that
            // are not part of the current group read the zero row of T1.
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int col = 2 * jl + h;
                const uint32_t row = col < ncur ? (uint32_t)(e1 + col) : zrow;
                const uint32_t ta = t1_lane + row * (uint32_t)(PS * 8);
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
                    acc_cur[mt] = fma(lds64(ta + 64 * mt), cacc[mt][h], acc_cur[mt]);
            }
            if (w0 & (1u << 21)) {
                // the current group ends: sum over the 4 lanes of a point, fold upwards
                const int nf = (w0 >> 22) & 15;
                const bool last = (w0 >> 26) & 1u;
                double low[MT];
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) low[mt] = quad_sum(acc_cur[mt]);
                if (nf == 0) {
                    // m == 2: single group, the program is over
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) a2[mt] = low[mt];
                } else if (nf == 1) {
                    const double g2 = gps[2 * p.gp_rows + (int)w1];
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) {
                        a2[mt] = fma(t2[mt], low[mt], a2[mt]);
                        if (!last) t2[mt] *= xs[pbase + 8 * mt] - g2;
                    }
                } else {
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) {
                        low[mt] = fma(t2[mt], low[mt], a2[mt]);
                        a2[mt] = 0.0;
                        t2[mt] = 1.0;
                    }
#pragma unroll 1
                    for (int u = 1; u < nf; ++u) {
                        double *ut = ups + (size_t)(u - 1) * PS + pbase;
                        const bool top = (u == nf - 1);
                        const double gu = gps[(2 + u) * p.gp_rows + (int)w1];
                        // the 4 lanes of a point compute and store identical values to the same
                        // words: all of them read first, the warp meets, then they store -- a lane
                        // that lags behind (thread 0 polls the ring in this loop) must not read
                        // what a faster lane has already restarted
                        double tvv[MT], avv[MT];
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
                            tvv[mt] = ut[8 * mt];
                            avv[mt] = fma(tvv[mt], low[mt], ut[up_stride + 8 * mt]);
                        }
                        __syncwarp();
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
                            if (!top) {
                                low[mt] = avv[mt];
                                ut[up_stride + 8 * mt] = 0.0;
                                ut[8 * mt] = 1.0;
                            } else {
                                ut[up_stride + 8 * mt] = avv[mt];
                                if (!last) ut[8 * mt] = tvv[mt] * (xs[(size_t)u * PS + pbase + 8 * mt] - gu);
                            }
                        }
                        __syncwarp();
                    }
                }
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) acc_cur[mt] = 0.0;
                if (nnext) {
                    // head of the next group in the same tile: its dimension-1 exponents start at 0
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int col = 2 * jl + h;
                        const uint32_t row = (col >= ncur && col < ncur + nnext) ? (uint32_t)(col - ncur) : zrow;
                        const uint32_t ta = t1_lane + row * (uint32_t)(PS * 8);
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt)
                            acc_cur[mt] = fma(lds64(ta + 64 * mt), cacc[mt][h], acc_cur[mt]);
                    }
                }
            }
        }
        mbar_arrive(&empty[s]);
    }

    // ---- result: every point is held by 4 lanes, lane jl == 0 writes ---------------------------
    if (jl == 0) {
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
            if (ptq[mt] < p.P) {
                const double res = (U <= 1) ? a2[mt] : ups[up_stride + (size_t)(U - 2) * PS + pbase + 8 * mt];
                p.out[ptq[mt] * p.ld_out + q] = res;
            }
        }
    }
}

template <int KS0, int NW, int MT, int MINB>
int launch_tmma(const TmmaParams &prm, int k, size_t smem, cudaStream_t stream)
{
    auto fn = &eval_tmma_kernel<KS0, NW, MT, MINB>;
    MPX_CUDA(mpx_allow_max_smem(fn));
    const int64_t ptiles = (prm.P + 8 * MT * NW - 1) / (8 * MT * NW);
    if (ptiles > 0x7fffffffLL || k > 65535)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "too many points or polynomials for one launch");
    fn<<<dim3((unsigned)ptiles, (unsigned)k, 1), 32 * NW, smem, stream>>>(prm);
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}

size_t tmma_smem_bytes(const mpx_plan *pl, int pts)
{
    const int ps = pts + 2;
    const int U = pl->m - 2;
    size_t doubles = (size_t)((pl->m * pl->gp_rows + 1) & ~1) + (size_t)(pl->max_exp[1] + 2) * ps +
                     (size_t)(U > 0 ? U : 0) * ps + (size_t)2 * (U > 1 ? U - 1 : 0) * ps;
    return 128 + (size_t)T_STAGES * T_STAGE_BYTES + doubles * 8;
}

}  // namespace

namespace {
int tmma_done(int rc, MpxImage *img, cudaStream_t stream)
{
    return rc != MPX_OK ? rc : mpx_image_release(img, stream);
}
}  // namespace

bool mpx_tmma_eligible(const mpx_plan *pl, int k)
{
    (void)k;
    return pl->tprog.usable && pl->d_tdesc != nullptr && tmma_smem_bytes(pl, 256) <= (size_t)227 * 1024;
}

size_t mpx_tmma_scratch_bytes(const mpx_plan *pl, int k)
{
    return (size_t)pl->tprog.n_chunks * MPX_T_COEF_CAP * (size_t)k * sizeof(double);
}

int mpx_launch_eval_tmma(mpx_plan *pl, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                         const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                         cudaStream_t stream)
{
    const int64_t n_slots = (int64_t)pl->tprog.n_chunks * MPX_T_COEF_CAP;
    if (coeffs) {
        int rc = mpx_image_acquire(img, mpx_tmma_scratch_bytes(pl, k), stream);
        if (rc != MPX_OK) return rc;
        const int64_t total = n_slots * k;
        pack_tile_coeffs_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
            coeffs, ld_c, k, pl->d_tslot_src, n_slots, img->ptr);
        MPX_CUDA(cudaGetLastError());
    }
    if (P == 0) return coeffs ? mpx_image_release(img, stream) : MPX_OK;
    TmmaParams prm;
    prm.xx = xx; prm.ld_xx = ld_xx; prm.P = P; prm.out = out; prm.ld_out = ld_out;
    prm.desc = pl->d_tdesc; prm.coef = img->ptr; prm.gp = pl->d_gp;
    prm.n_chunks = pl->tprog.n_chunks; prm.m = pl->m; prm.gp_rows = pl->gp_rows;
    for (int i = 0; i < 8; ++i) prm.max_exp[i] = i < pl->m ? pl->max_exp[i] : 0;
    const int ks = pl->tprog.ks_max;
#ifdef MPX_TUNING
    static const int env_mt = getenv("MPX_TMMA_MT") ? atoi(getenv("MPX_TMMA_MT")) : 0;   // experiment knob
#else
    constexpr int env_mt = 0;
#endif
    const bool mt2 = env_mt ? env_mt == 2 : ks > 5;
    const size_t smem = tmma_smem_bytes(pl, mt2 ? 128 : 256);
    if (mt2) {
        if (ks <= 3) return tmma_done(launch_tmma<3, 8, 2, 3>(prm, k, smem, stream), img, stream);
        if (ks <= 5) return tmma_done(launch_tmma<5, 8, 2, 3>(prm, k, smem, stream), img, stream);
        if (ks <= 9) return tmma_done(launch_tmma<9, 8, 2, 2>(prm, k, smem, stream), img, stream);
        if (ks <= 17) return tmma_done(launch_tmma<17, 8, 2, 1>(prm, k, smem, stream), img, stream);
    } else {
        if (ks <= 3) return tmma_done(launch_tmma<3, 8, 4, 2>(prm, k, smem, stream), img, stream);
        if (ks <= 5) return tmma_done(launch_tmma<5, 8, 4, 2>(prm, k, smem, stream), img, stream);
        if (ks <= 9) return tmma_done(launch_tmma<9, 8, 4, 1>(prm, k, smem, stream), img, stream);
        if (ks <= 17) return tmma_done(launch_tmma<17, 8, 4, 1>(prm, k, smem, stream), img, stream);
    }
    return mpx_fail(MPX_ERR_UNSUPPORTED, "dimension-0 table too long for the tensor-core tree kernel");
}
