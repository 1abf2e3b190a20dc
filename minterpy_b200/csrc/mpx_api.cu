// extern "C" entry points: argument checks, device selection, host-buffer variants.
#include <algorithm>

#include "mpx_common.h"

namespace {
int check_plan(const mpx_plan *pl)
{
    if (!pl) return mpx_fail(MPX_ERR_ARG, "plan is null");
    return MPX_OK;
}
}  // namespace

extern "C" int mpx_eval(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx,
                        const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                        void *stream)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (P < 0 || k < 1) return mpx_fail(MPX_ERR_ARG, "P must be >= 0 and k >= 1");
    if (P == 0) return MPX_OK;
    if (!xx || !coeffs || !out) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld_xx < pl->m || ld_c < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    std::lock_guard<std::mutex> lock(pl->mu);
    return mpx_launch_eval(pl, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, (cudaStream_t)stream);
}

extern "C" int mpx_eval_deriv(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx,
                              const double *coeffs, int k, int64_t ld_c, const int32_t *orders_h,
                              double *out, int64_t ld_out, void *stream)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (!orders_h) return mpx_fail(MPX_ERR_ARG, "orders_h is null");
    int orders[MPX_MAX_M];
    bool any = false;
    for (int i = 0; i < pl->m; ++i) {
        if (orders_h[i] < 0) return mpx_fail(MPX_ERR_ARG, "derivative orders must be >= 0");
        orders[i] = orders_h[i];
        any = any || orders[i] > 0;
    }
    // order 0 everywhere is the polynomial itself (the reference builds the plain tables then)
    if (!any) return mpx_eval(pl, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, stream);
    if (P < 0 || k < 1) return mpx_fail(MPX_ERR_ARG, "P must be >= 0 and k >= 1");
    if (P == 0) return MPX_OK;
    if (!xx || !coeffs || !out) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld_xx < pl->m || ld_c < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    if (!mpx_mma_fits(pl, k))
        return mpx_fail(MPX_ERR_UNSUPPORTED,
                        "derivative evaluation runs on the shared-memory-table kernel: m <= 8 and "
                        "m * (largest exponent + 1) * 68 doubles of tables must fit shared memory");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    std::lock_guard<std::mutex> lock(pl->mu);
    return mpx_launch_eval_mma(pl, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, (cudaStream_t)stream,
                               orders);
}

extern "C" int mpx_monomials(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx,
                             double *out, int64_t ld_out, int triangular, void *stream)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (P < 0) return mpx_fail(MPX_ERR_ARG, "P must be >= 0");
    if (P == 0) return MPX_OK;
    if (!xx || !out) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld_xx < pl->m || ld_out < pl->N) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    return mpx_launch_monomials(pl, xx, P, ld_xx, out, ld_out, triangular, (cudaStream_t)stream);
}

extern "C" int mpx_dds(mpx_plan *pl, double *vals, int k, int64_t ld, void *stream)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (k < 1) return mpx_fail(MPX_ERR_ARG, "k must be >= 1");
    if (!vals) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    return mpx_launch_dds(pl, vals, k, ld, (cudaStream_t)stream);
}

// ---- host-buffer variants ----------------------------------------------------------------
namespace {

int grow(double **buf, size_t *cap, size_t bytes)
{
    if (bytes <= *cap) return MPX_OK;
    if (*buf) MPX_CUDA(cudaFree(*buf));
    *buf = nullptr;
    *cap = 0;
    MPX_CUDA(cudaMalloc((void **)buf, bytes));
    *cap = bytes;
    return MPX_OK;
}

// row-major (rows x cols) copy between a pitched and a dense buffer; a dense source/target
// pair degenerates to one linear copy
cudaError_t copy_rows(void *dst, size_t dst_ld, const void *src, size_t src_ld, size_t cols,
                      size_t rows, cudaMemcpyKind kind, cudaStream_t st)
{
    if (dst_ld == cols && src_ld == cols)
        return cudaMemcpyAsync(dst, src, rows * cols * 8, kind, st);
    return cudaMemcpy2DAsync(dst, dst_ld * 8, src, src_ld * 8, cols * 8, rows, kind, st);
}

}  // namespace

// Evaluation with host buffers.  The coefficients are uploaded and packed once; the points go
// through two device blocks on two streams, so that the upload of block b+1 and the download
// of block b-1 overlap the evaluation of block b.  Staging buffers and streams belong to the
// plan and are reused by later calls.
extern "C" int mpx_eval_host(mpx_plan *pl, const double *xx_h, int64_t P, int64_t ld_xx,
                             const double *coeffs_h, int k, int64_t ld_c, double *out_h,
                             int64_t ld_out)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (P < 0 || k < 1) return mpx_fail(MPX_ERR_ARG, "P must be >= 0 and k >= 1");
    if (P == 0) return MPX_OK;
    if (!xx_h || !coeffs_h || !out_h) return mpx_fail(MPX_ERR_ARG, "null host pointer");
    const int m = pl->m;
    if (ld_xx < m || ld_c < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    std::lock_guard<std::mutex> lock(pl->mu);

    // block size: ~64 MiB of points+values per block, at least 64 Ki points
    int64_t block = ((int64_t)64 << 20) / (8 * (m + k));
    block = std::max<int64_t>(block, (int64_t)1 << 16);
    block = std::min(block, P);
    const int nbuf = (P > block) ? 2 : 1;
    for (int b = 0; b < 2; ++b)
        if (!pl->h_stream[b]) MPX_CUDA(cudaStreamCreateWithFlags(&pl->h_stream[b], cudaStreamNonBlocking));
    if (!pl->h_event) MPX_CUDA(cudaEventCreateWithFlags(&pl->h_event, cudaEventDisableTiming));
    if ((rc = grow(&pl->h_c, &pl->h_c_cap, (size_t)pl->N * k * 8)) != MPX_OK) return rc;
    for (int b = 0; b < nbuf; ++b) {
        if ((rc = grow(&pl->h_x[b], &pl->h_x_cap[b], (size_t)block * m * 8)) != MPX_OK) return rc;
        if ((rc = grow(&pl->h_o[b], &pl->h_o_cap[b], (size_t)block * k * 8)) != MPX_OK) return rc;
    }

    cudaStream_t s0 = pl->h_stream[0];
    MPX_CUDA(copy_rows(pl->h_c, k, coeffs_h, ld_c, k, pl->N, cudaMemcpyHostToDevice, s0));
    if ((rc = mpx_launch_eval(pl, nullptr, 0, m, pl->h_c, k, k, nullptr, k, s0)) != MPX_OK) return rc;
    MPX_CUDA(cudaEventRecord(pl->h_event, s0));
    MPX_CUDA(cudaStreamWaitEvent(pl->h_stream[1], pl->h_event, 0));

    int status = MPX_OK;
    const int64_t nblk = (P + block - 1) / block;
    for (int64_t b = 0; b < nblk && status == MPX_OK; ++b) {
        const int s = (int)(b % nbuf);
        cudaStream_t st = pl->h_stream[s];
        const int64_t p0 = b * block, np = std::min(block, P - p0);
        cudaError_t e = copy_rows(pl->h_x[s], m, xx_h + p0 * ld_xx, ld_xx, m, np,
                                  cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) { status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e)); break; }
        status = mpx_launch_eval(pl, pl->h_x[s], np, m, nullptr, k, k, pl->h_o[s], k, st);
        if (status != MPX_OK) break;
        e = copy_rows(out_h + p0 * ld_out, ld_out, pl->h_o[s], k, k, np, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) { status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e)); break; }
    }
    for (int b = 0; b < 2; ++b) {
        cudaError_t e = cudaStreamSynchronize(pl->h_stream[b]);
        if (e != cudaSuccess && status == MPX_OK) status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e));
    }
    return status;
}

extern "C" int mpx_dds_host(mpx_plan *pl, const double *in_h, int k, int64_t ld_in, double *out_h,
                            int64_t ld_out)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (k < 1) return mpx_fail(MPX_ERR_ARG, "k must be >= 1");
    if (!in_h || !out_h) return mpx_fail(MPX_ERR_ARG, "null host pointer");
    if (ld_in < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    if (!pl->complete) return mpx_launch_dds(pl, nullptr, k, k, nullptr);   // reports INCOMPLETE
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    double *d_v = nullptr;
    MPX_CUDA(cudaMalloc((void **)&d_v, (size_t)pl->N * k * sizeof(double)));
    int status = MPX_OK;
    cudaError_t e = cudaMemcpy2D(d_v, (size_t)k * 8, in_h, (size_t)ld_in * 8, (size_t)k * 8, pl->N,
                                 cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        status = mpx_launch_dds(pl, d_v, k, k, nullptr);
        if (status == MPX_OK)
            e = cudaMemcpy2D(out_h, (size_t)ld_out * 8, d_v, (size_t)k * 8, (size_t)k * 8, pl->N,
                             cudaMemcpyDeviceToHost);
    }
    if (e != cudaSuccess) status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e));
    cudaFree(d_v);
    return status;
}
