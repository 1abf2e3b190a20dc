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
// Evaluation: points are processed in blocks; block b+1 is uploaded on one stream while
// block b is evaluated and block b-1 is downloaded (two device buffers, two streams).
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

    const int64_t block = std::min<int64_t>(P, (int64_t)1 << 22);   // 4 Mi points per block
    const int nbuf = (P > block) ? 2 : 1;
    double *d_c = nullptr, *d_x[2] = {nullptr, nullptr}, *d_o[2] = {nullptr, nullptr};
    cudaStream_t st[2] = {nullptr, nullptr};
    int status = MPX_OK;
    auto cleanup = [&]() {
        cudaFree(d_c);
        for (int b = 0; b < 2; ++b) {
            cudaFree(d_x[b]);
            cudaFree(d_o[b]);
            if (st[b]) cudaStreamDestroy(st[b]);
        }
    };
#define MPX_TRY(expr)                                                                          \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            status = mpx_fail(MPX_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
            cleanup();                                                                         \
            return status;                                                                     \
        }                                                                                      \
    } while (0)
    MPX_TRY(cudaMalloc((void **)&d_c, (size_t)pl->N * k * sizeof(double)));
    MPX_TRY(cudaMemcpy2D(d_c, (size_t)k * 8, coeffs_h, (size_t)ld_c * 8, (size_t)k * 8, pl->N,
                         cudaMemcpyHostToDevice));
    for (int b = 0; b < nbuf; ++b) {
        MPX_TRY(cudaMalloc((void **)&d_x[b], (size_t)block * m * sizeof(double)));
        MPX_TRY(cudaMalloc((void **)&d_o[b], (size_t)block * k * sizeof(double)));
        MPX_TRY(cudaStreamCreateWithFlags(&st[b], cudaStreamNonBlocking));
    }
    {
        // one stream at a time may use the plan's scratch: pack once by evaluating block 0
        // first; later blocks repack the same coefficients (cheap), serialised by an event
        std::lock_guard<std::mutex> lock(pl->mu);
        cudaEvent_t ev_prev = nullptr;
        int64_t nblk = (P + block - 1) / block;
        for (int64_t b = 0; b < nblk && status == MPX_OK; ++b) {
            const int s = (int)(b % nbuf);
            const int64_t p0 = b * block, np = std::min(block, P - p0);
            cudaError_t e = cudaMemcpy2DAsync(d_x[s], (size_t)m * 8, xx_h + p0 * ld_xx,
                                              (size_t)ld_xx * 8, (size_t)m * 8, np,
                                              cudaMemcpyHostToDevice, st[s]);
            if (e == cudaSuccess && ev_prev) e = cudaStreamWaitEvent(st[s], ev_prev, 0);
            if (e != cudaSuccess) { status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e)); break; }
            status = mpx_launch_eval(pl, d_x[s], np, m, d_c, k, k, d_o[s], k, st[s]);
            if (status != MPX_OK) break;
            if (ev_prev) cudaEventDestroy(ev_prev);
            ev_prev = nullptr;
            e = cudaEventCreateWithFlags(&ev_prev, cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventRecord(ev_prev, st[s]);
            if (e == cudaSuccess)
                e = cudaMemcpy2DAsync(out_h + p0 * ld_out, (size_t)ld_out * 8, d_o[s], (size_t)k * 8,
                                      (size_t)k * 8, np, cudaMemcpyDeviceToHost, st[s]);
            if (e != cudaSuccess) { status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e)); break; }
        }
        if (ev_prev) cudaEventDestroy(ev_prev);
        for (int b = 0; b < nbuf; ++b) {
            cudaError_t e = cudaStreamSynchronize(st[b]);
            if (e != cudaSuccess && status == MPX_OK) status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e));
        }
    }
    cleanup();
    return status;
#undef MPX_TRY
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
