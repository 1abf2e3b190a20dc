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
#include "mpx_common.h"

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

}  // namespace

int mpx_launch_dds(mpx_plan *pl, double *vals, int k, int64_t ld, cudaStream_t stream)
{
    if (!pl->complete)
        return mpx_fail(MPX_ERR_INCOMPLETE,
                        "trying to use the divided difference scheme with incomplete multi "
                        "indices, but DDS only works for complete multi indices (without 'holes').");
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
