// K4: dense operator apply  out = T @ c   (FP64).
//   replaces MatrixOperator.__matmul__, src/minterpy/schemes/matrix_operator.py:13-22.
// T is (rows x cols); for the transformations of this path it is N x N, read once from HBM
// (8 N^2 bytes) and that read bounds the kernel for every k the library meets.  A CTA owns
// 32 rows; a warp owns 4 of them and sweeps the columns, each lane keeping KT partial sums per
// row, reduced by shuffles at the end.  The coefficient rows a span of columns needs are staged
// transposed in shared memory by the whole CTA (one barrier per 64 KB of matrix), so the inner
// loop reads them with conflict-free 16-byte loads whatever k and ld_c are.  Round 2: this FMA
// form serves one or two coefficient columns; from three on the tensor-core kernel below reads T
// once for up to 32 columns (k = 8: 1.07 ms against 1.55 ms here at N = 24 809).
#include "mpx_common.h"

namespace {

constexpr int ROWS_PER_WARP = 4;
constexpr int WARPS = 8;
constexpr int NTHREADS = WARPS * 32;
constexpr int UNR = 4;            // column strips of 64 in flight per row (double2 loads)
constexpr int SPAN = 64 * UNR;    // columns per staged span
constexpr int CS = SPAN + 4;      // pitch of a staged coefficient row (two columns of overlap)

// lanes sweep the columns two at a time (16-byte loads), UNR strips per row are issued before
// any of them is used: 4 rows x 4 strips x 16 B = 256 B in flight per lane.
// ODD: the leading dimension is odd (the usual case, N x N with N odd), so rows start
// alternately 0 and 8 bytes past a 16-byte boundary.  A warp's first row is a multiple of 4,
// hence its rows 0 and 2 are aligned and rows 1 and 3 are not: those load their pairs one
// column further right (columns j+1, j+2 instead of j, j+1), their column 0 is added by lane 0.
// KT = coefficient columns per pass (grid.y covers k in steps of KT).
template <bool ODD, int KT, bool LOWER>
__global__ void __launch_bounds__(NTHREADS)
dense_apply_kernel(const double *__restrict__ T, int64_t rows, int64_t cols_all, int64_t ld_t,
                   const double *__restrict__ c, int k, int64_t ld_c, double *__restrict__ out,
                   int64_t ld_out)
{
    constexpr bool lower = LOWER;
    __shared__ __align__(16) double cs[2][KT][CS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // lower-triangular operator: the heavy row blocks (bottom of the matrix) are scheduled
    // first, and a CTA sweeps the columns up to its last row only -- what lies to the right is
    // zero by contract and never read (the few stored zeros inside the block's last columns are)
    const int64_t cta = lower ? (int64_t)gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const int64_t cta_end = (cta + 1) * WARPS * ROWS_PER_WARP;
    const int64_t cols = (lower && cta_end < cols_all) ? cta_end : cols_all;
    const int64_t r0 = (cta * WARPS + warp) * ROWS_PER_WARP;
    const int q0 = blockIdx.y * KT;
    const bool active = r0 < rows;
    // the last warp may own fewer than 4 real rows; its stand-in rows do not follow the
    // alignment pattern, so it takes the plain loop at the bottom for all columns
    const bool full = r0 + ROWS_PER_WARP <= rows;
    double acc[ROWS_PER_WARP][KT];
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
        for (int q = 0; q < KT; ++q) acc[i][q] = 0.0;
    const double *row[ROWS_PER_WARP];
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i) row[i] = T + ((r0 + i < rows) ? r0 + i : rows - 1) * ld_t;

    constexpr int STAGE = (SPAN + 2) * KT;                    // staged doubles per span
    constexpr int NS = (STAGE + NTHREADS - 1) / NTHREADS;     // of them per thread
    const int64_t n_spans = (cols - (ODD ? 1 : 0)) / SPAN;    // spans with SPAN (+1) columns left
    double creg[NS];
    auto stage_load = [&](int64_t sp) {
#pragma unroll
        for (int n = 0; n < NS; ++n) {
            const int idx = tid + n * NTHREADS;
            const int jj = idx / KT, q = idx % KT;
            const int64_t col = sp * SPAN + jj;
            creg[n] = (idx < STAGE && col < cols && q0 + q < k) ? c[col * ld_c + q0 + q] : 0.0;
        }
    };
    auto stage_store = [&](int buf) {
#pragma unroll
        for (int n = 0; n < NS; ++n) {
            const int idx = tid + n * NTHREADS;
            if (idx < STAGE) cs[buf][idx % KT][idx / KT] = creg[n];
        }
    };
    if (n_spans > 0) {
        stage_load(0);
        stage_store(0);
    }
    __syncthreads();
    for (int64_t sp = 0; sp < n_spans; ++sp) {
        const int buf = (int)(sp & 1);
        const int64_t j0 = sp * SPAN;
        if (sp + 1 < n_spans) stage_load(sp + 1);
        if (full) {
            double2 t[ROWS_PER_WARP][UNR];
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
                for (int u = 0; u < UNR; ++u)
                    t[i][u] = __ldcs(reinterpret_cast<const double2 *>(
                        row[i] + j0 + 64 * u + 2 * lane + ((ODD && (i & 1)) ? 1 : 0)));
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int jj = 64 * u + 2 * lane;
#pragma unroll
                for (int q = 0; q < KT; ++q) {
                    const double2 c01 = *reinterpret_cast<const double2 *>(&cs[buf][q][jj]);
                    const double c2 = ODD ? cs[buf][q][jj + 2] : 0.0;
#pragma unroll
                    for (int i = 0; i < ROWS_PER_WARP; ++i) {
                        const bool shifted = ODD && (i & 1);
                        acc[i][q] = fma(t[i][u].y, shifted ? c2 : c01.y,
                                        fma(t[i][u].x, shifted ? c01.y : c01.x, acc[i][q]));
                    }
                }
            }
        }
        if (sp + 1 < n_spans) stage_store(buf ^ 1);
        __syncthreads();
    }
    if (!active) return;
    int64_t j0 = 0;
    if (full) {
        j0 = n_spans * SPAN;
        if (ODD && lane == 0) {                  // column 0 of the shifted rows
#pragma unroll
            for (int q = 0; q < KT; ++q) {
                const double cv = (q0 + q < k) ? c[q0 + q] : 0.0;
                acc[1][q] = fma(row[1][0], cv, acc[1][q]);
                acc[3][q] = fma(row[3][0], cv, acc[3][q]);
            }
        }
    }
    // remaining columns: [j0, cols) for the unshifted rows, [j0 + 1, cols) for the shifted ones
    for (int64_t j = j0 + lane; j < cols; j += 32) {
        double cv[KT];
#pragma unroll
        for (int q = 0; q < KT; ++q) cv[q] = (q0 + q < k) ? c[j * ld_c + q0 + q] : 0.0;
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i) {
            if (ODD && full && (i & 1) && j == j0) continue;
            const double t = row[i][j];
#pragma unroll
            for (int q = 0; q < KT; ++q) acc[i][q] = fma(t, cv[q], acc[i][q]);
        }
    }
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
        for (int q = 0; q < KT; ++q) {
            double v = acc[i][q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && r0 + i < rows && q0 + q < k) out[(r0 + i) * ld_out + q0 + q] = v;
        }
}

// the matrix itself is not 16-byte aligned (a view that starts at an odd column): 8-byte
// loads, 4 strips of 32 columns in flight per row, 4 coefficient columns per pass
constexpr int KT_PLAIN = 4;
__global__ void __launch_bounds__(NTHREADS)
dense_apply_plain_kernel(const double *__restrict__ T, int64_t rows, int64_t cols_all, int64_t ld_t,
                         const double *__restrict__ c, int k, int64_t ld_c,
                         double *__restrict__ out, int64_t ld_out, int lower)
{
    constexpr int KT = KT_PLAIN;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t cta = lower ? (int64_t)gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const int64_t cta_end = (cta + 1) * WARPS * ROWS_PER_WARP;
    const int64_t cols = (lower && cta_end < cols_all) ? cta_end : cols_all;
    const int64_t r0 = (cta * WARPS + warp) * ROWS_PER_WARP;
    const int q0 = blockIdx.y * KT;
    if (r0 >= rows) return;
    double acc[ROWS_PER_WARP][KT];
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
        for (int q = 0; q < KT; ++q) acc[i][q] = 0.0;
    const double *row[ROWS_PER_WARP];
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i) row[i] = T + ((r0 + i < rows) ? r0 + i : rows - 1) * ld_t;
    int64_t j0 = 0;
    for (; j0 + 32 * UNR <= cols; j0 += 32 * UNR) {
        double t[ROWS_PER_WARP][UNR];
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
            for (int u = 0; u < UNR; ++u) t[i][u] = __ldcs(row[i] + j0 + 32 * u + lane);
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int64_t j = j0 + 32 * u + lane;
#pragma unroll
            for (int q = 0; q < KT; ++q) {
                const double cv = (q0 + q < k) ? c[j * ld_c + q0 + q] : 0.0;
#pragma unroll
                for (int i = 0; i < ROWS_PER_WARP; ++i) acc[i][q] = fma(t[i][u], cv, acc[i][q]);
            }
        }
    }
    for (int64_t j = j0 + lane; j < cols; j += 32) {
#pragma unroll
        for (int q = 0; q < KT; ++q) {
            const double cv = (q0 + q < k) ? c[j * ld_c + q0 + q] : 0.0;
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) acc[i][q] = fma(row[i][j], cv, acc[i][q]);
        }
    }
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
        for (int q = 0; q < KT; ++q) {
            double v = acc[i][q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && r0 + i < rows && q0 + q < k) out[(r0 + i) * ld_out + q0 + q] = v;
        }
}

template <int KT>
void launch_vec(bool odd, dim3 grid, cudaStream_t st, const double *T, int64_t rows, int64_t cols,
                int64_t ld_t, const double *c, int k, int64_t ld_c, double *out, int64_t ld_out,
                int lower)
{
    if (odd && lower)
        dense_apply_kernel<true, KT, true><<<grid, NTHREADS, 0, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out);
    else if (odd)
        dense_apply_kernel<true, KT, false><<<grid, NTHREADS, 0, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out);
    else if (lower)
        dense_apply_kernel<false, KT, true><<<grid, NTHREADS, 0, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out);
    else
        dense_apply_kernel<false, KT, false><<<grid, NTHREADS, 0, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out);
}

// ---- many coefficient columns: one pass over T on the FP64 tensor cores ------------------------
// out (rows x k) = T (rows x cols) c (cols x k) as DMMA m8n8k4 (mma.sync; sm_100a has no tcgen05
// kind for f64).  A warp owns 8 rows (one m-tile) and 8 NT coefficient columns (NT accumulator
// tiles, 2 NT doubles per lane) and sweeps the columns of T once, 16 at a time: lane (g, t) =
// (lane / 4, lane % 4) loads T[row g][16 s + 4 t .. + 3] -- 32 contiguous bytes, a whole 128-byte
// line per row and quarter-warp -- and uses its j-th element as the A-fragment entry (row g,
// k-index t) of the j-th of four k-steps.  The contraction index is only PERMUTED by that (step j
// covers columns 16 s + 4 t + j), and the B fragments follow: lane (t, n = g) reads
// c[16 s + 4 t + j][8 nt + n] from the span of coefficient rows the CTA staged in shared memory
// (pitch 8 NT + 1 doubles: conflict free).  T is read exactly once for up to 32 columns; the
// kernel stays bound by that read up to k = 16 and by the tensor pipe beyond.
constexpr int MW = 4;              // warps (m-tiles) per CTA
constexpr int MSPAN = 128;         // columns of T per staged span of coefficient rows
constexpr int MSUP = MSPAN / 16;   // super-steps of 16 columns per span

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

template <int NT, bool VEC>
__global__ void __launch_bounds__(MW * 32)
dense_mma_kernel(const double *__restrict__ T, int64_t rows, int64_t cols_all, int64_t ld_t,
                 const double *__restrict__ c, int k, int64_t ld_c, double *__restrict__ out,
                 int64_t ld_out, int lower)
{
    constexpr int KP = 8 * NT + 1;
    extern __shared__ __align__(16) double msm[];                 // [2][MSPAN][KP]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int64_t cta = lower ? (int64_t)gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const int64_t cta_end = (cta + 1) * MW * 8;
    const int64_t cols = (lower && cta_end < cols_all) ? cta_end : cols_all;
    const int64_t r = (cta * MW + warp) * 8 + g;                  // this lane's row of T
    const double *row = T + (r < rows ? r : rows - 1) * ld_t;
    const int q0 = blockIdx.y * (8 * NT);
    double acc[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) acc[nt][0] = acc[nt][1] = 0.0;

    constexpr int STAGE = MSPAN * 8 * NT;
    constexpr int NS = STAGE / (MW * 32);
    double creg[NS];
    auto stage_load = [&](int64_t j0) {
#pragma unroll
        for (int n = 0; n < NS; ++n) {
            const int idx = tid + n * (MW * 32);
            const int jj = idx / (8 * NT), q = idx % (8 * NT);
            creg[n] = (j0 + jj < cols && q0 + q < k) ? c[(j0 + jj) * ld_c + q0 + q] : 0.0;
        }
    };
    auto stage_store = [&](int buf) {
#pragma unroll
        for (int n = 0; n < NS; ++n) {
            const int idx = tid + n * (MW * 32);
            msm[((size_t)buf * MSPAN + idx / (8 * NT)) * KP + idx % (8 * NT)] = creg[n];
        }
    };
    const int64_t n_spans = (cols + MSPAN - 1) / MSPAN;
    stage_load(0);
    stage_store(0);
    __syncthreads();
    for (int64_t sp = 0; sp < n_spans; ++sp) {
        const int buf = (int)(sp & 1);
        const int64_t j0 = sp * MSPAN;
        if (sp + 1 < n_spans) stage_load(j0 + MSPAN);
        // all A values of the span first: 8 super-steps x 32 bytes in flight per lane
        double a[MSUP][4];
        if (j0 + MSPAN <= cols) {
#pragma unroll
            for (int s = 0; s < MSUP; ++s) {
                const double *p = row + j0 + 16 * s + 4 * t;
                if constexpr (VEC) {
                    const double2 v0 = __ldcs(reinterpret_cast<const double2 *>(p));
                    const double2 v1 = __ldcs(reinterpret_cast<const double2 *>(p + 2));
                    a[s][0] = v0.x; a[s][1] = v0.y; a[s][2] = v1.x; a[s][3] = v1.y;
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) a[s][j] = __ldcs(p + j);
                }
            }
        } else {                                   // last span: columns past the end are zeros
#pragma unroll
            for (int s = 0; s < MSUP; ++s)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int64_t col = j0 + 16 * s + 4 * t + j;
                    a[s][j] = col < cols ? row[col] : 0.0;
                }
        }
        const double *cb = msm + (size_t)buf * MSPAN * KP;
#pragma unroll
        for (int s = 0; s < MSUP; ++s)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const double *bp = cb + (16 * s + 4 * t + j) * KP + g;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(acc[nt][0], acc[nt][1], a[s][j], bp[8 * nt]);
            }
        if (sp + 1 < n_spans) stage_store(buf ^ 1);
        __syncthreads();
    }
    if (r < rows) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int q = q0 + 8 * nt + 2 * t;
            if (q < k) out[r * ld_out + q] = acc[nt][0];
            if (q + 1 < k) out[r * ld_out + q + 1] = acc[nt][1];
        }
    }
}

template <int NT>
int launch_mma(const double *T, int64_t rows, int64_t cols, int64_t ld_t, const double *c, int k,
               int64_t ld_c, double *out, int64_t ld_out, int lower, cudaStream_t st)
{
    const size_t smem = (size_t)2 * MSPAN * (8 * NT + 1) * sizeof(double);
    dim3 grid((unsigned)((rows + MW * 8 - 1) / (MW * 8)), (unsigned)((k + 8 * NT - 1) / (8 * NT)), 1);
    const bool vec = (reinterpret_cast<uintptr_t>(T) & 15) == 0 && (ld_t & 1) == 0;
    if (vec) {
        MPX_CUDA(mpx_allow_max_smem(dense_mma_kernel<NT, true>));
        dense_mma_kernel<NT, true><<<grid, MW * 32, smem, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower);
    } else {
        MPX_CUDA(mpx_allow_max_smem(dense_mma_kernel<NT, false>));
        dense_mma_kernel<NT, false><<<grid, MW * 32, smem, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower);
    }
    return MPX_OK;
}

}  // namespace

extern "C" int mpx_apply_dense_ex(int device, const double *T, int64_t rows, int64_t cols,
                                  int64_t ld_t, const double *c, int k, int64_t ld_c, double *out,
                                  int64_t ld_out, int flags, void *stream)
{
    if (rows < 0 || cols < 0 || k < 1) return mpx_fail(MPX_ERR_ARG, "bad sizes");
    if (flags & ~MPX_DENSE_LOWER) return mpx_fail(MPX_ERR_ARG, "unknown flag");
    if (rows == 0) return MPX_OK;
    if (!T || !c || !out) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld_t < cols || ld_c < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select CUDA device");
    const int lower = (flags & MPX_DENSE_LOWER) ? 1 : 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (k > 2) {
        // one pass over T for up to 32 columns on the tensor cores (8 or 16 per pass up to
        // k = 16: smaller tiles, more CTAs per SM)
        int rc = k <= 8    ? launch_mma<1>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower, st)
                 : k <= 16 ? launch_mma<2>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower, st)
                           : launch_mma<4>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower, st);
        if (rc != MPX_OK) return rc;
        MPX_CUDA(cudaGetLastError());
        return MPX_OK;
    }
    const int64_t rows_per_cta = (int64_t)WARPS * ROWS_PER_WARP;
    const unsigned gx = (unsigned)((rows + rows_per_cta - 1) / rows_per_cta);
    const bool aligned = (reinterpret_cast<uintptr_t>(T) & 15) == 0;
    if (!aligned) {
        dim3 grid(gx, (unsigned)((k + KT_PLAIN - 1) / KT_PLAIN), 1);
        dense_apply_plain_kernel<<<grid, NTHREADS, 0, st>>>(T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower);
    } else {
        const bool odd = (ld_t & 1) != 0;
        dim3 grid(gx, 1, 1);                                 // k = 1 or 2: one pass
        if (k == 1) launch_vec<1>(odd, grid, st, T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower);
        else launch_vec<2>(odd, grid, st, T, rows, cols, ld_t, c, k, ld_c, out, ld_out, lower);
    }
    MPX_CUDA(cudaGetLastError());
    return MPX_OK;
}

extern "C" int mpx_apply_dense(int device, const double *T, int64_t rows, int64_t cols,
                               int64_t ld_t, const double *c, int k, int64_t ld_c, double *out,
                               int64_t ld_out, void *stream)
{
    return mpx_apply_dense_ex(device, T, rows, cols, ld_t, c, k, ld_c, out, ld_out, 0, stream);
}
