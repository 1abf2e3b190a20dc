// extern "C" entry points: argument checks, device selection, host-buffer variants.
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

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
    return mpx_launch_eval(pl, &pl->image, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out, (cudaStream_t)stream);
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
    if (pl->tree_ok && !mpx_mma_eligible(pl, k)) {
        // few polynomials: the tree kernel with differentiated tables (mpx_eval_deriv.cu)
        DeviceGuard dgt(pl->device);
        if (!dgt.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
        std::lock_guard<std::mutex> lock(pl->mu);
        return mpx_launch_eval_tree_deriv(pl, &pl->image, &pl->deriv_tables, xx, P, ld_xx, coeffs, k,
                                          ld_c, orders, out, ld_out, (cudaStream_t)stream);
    }
    if (!mpx_mma_fits(pl, k))
        return mpx_fail(MPX_ERR_UNSUPPORTED,
                        "derivative evaluation runs on the shared-memory-table kernel: m <= 8 and "
                        "m * (largest exponent + 1) * 68 doubles of tables must fit shared memory");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    std::lock_guard<std::mutex> lock(pl->mu);
    return mpx_launch_eval_mma(pl, &pl->image, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out,
                               (cudaStream_t)stream, orders);
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
    return mpx_dds_ex(pl, vals, k, ld, 0, stream);
}

extern "C" int mpx_dds_ex(mpx_plan *pl, double *vals, int k, int64_t ld, int flags, void *stream)
{
    int rc = check_plan(pl);
    if (rc) return rc;
    if (k < 1) return mpx_fail(MPX_ERR_ARG, "k must be >= 1");
    if (!vals) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    if (flags & ~MPX_DDS_MANY_ZEROS) return mpx_fail(MPX_ERR_ARG, "unknown dds flag");
    return mpx_launch_dds(pl, vals, k, ld, (cudaStream_t)stream, flags);
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

int grow_pinned(double **buf, size_t *cap, size_t bytes)
{
    if (bytes <= *cap) return MPX_OK;
    if (*buf) MPX_CUDA(cudaFreeHost(*buf));
    *buf = nullptr;
    *cap = 0;
    MPX_CUDA(cudaHostAlloc((void **)buf, bytes, cudaHostAllocDefault));
    *cap = bytes;
    return MPX_OK;
}

// Is this host range ordinary pageable memory?  (Page-locked or registered memory can be handed
// to the copy engines directly; pageable memory is staged through the plan's page-locked blocks
// by host threads, which is several times faster than letting the driver stage it.)
bool is_pageable(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
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

// A few helper threads for repacking pageable arrays (one core moves ~10 GB/s, the PCIe link
// 55 GB/s).  They are started on first use and parked on a condition variable between jobs:
// starting threads per block costs more than copying a block.  The pool is never destroyed (the
// threads would have to be joined during process exit, when this library's state is already
// being torn down); one job at a time, guarded by `gate`.
class CopyPool {
public:
    static CopyPool &get()
    {
        static CopyPool *pool = new CopyPool();       // intentionally leaked
        return *pool;
    }
    // run fn(part) for part = 0..parts-1 on the helpers and the calling thread
    void run(int parts, const std::function<void(int)> &fn)
    {
        // one job at a time: callers feeding several GPUs from one process queue here (letting a
        // second caller copy its block alone instead was measured: 103 against 85 ms at 8 GPUs)
        std::lock_guard<std::mutex> one_job(gate);
        {
            std::lock_guard<std::mutex> lk(mu);
            job = &fn;
            next = 0;
            total = parts;
            pending = parts;
            ++generation;
        }
        wake.notify_all();
        work();
        std::unique_lock<std::mutex> lk(mu);
        done.wait(lk, [&] { return pending == 0; });
        job = nullptr;
    }
    int size() const { return (int)threads.size() + 1; }

private:
    CopyPool()
    {
        unsigned hw = std::thread::hardware_concurrency();
        const int helpers = (int)std::min(15u, hw > 1 ? hw - 1 : 0u);
        for (int t = 0; t < helpers; ++t) {
            threads.emplace_back([this] {
                uint64_t seen = 0;
                for (;;) {
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        wake.wait(lk, [&] { return generation != seen; });
                        seen = generation;
                    }
                    work();
                }
            });
            threads.back().detach();
        }
    }
    void work()
    {
        for (;;) {
            int part;
            const std::function<void(int)> *fn;
            {
                std::lock_guard<std::mutex> lk(mu);
                if (!job || next >= total) return;
                part = next++;
                fn = job;
            }
            (*fn)(part);
            std::lock_guard<std::mutex> lk(mu);
            if (--pending == 0) done.notify_all();
        }
    }
    std::mutex gate, mu;
    std::condition_variable wake, done;
    std::vector<std::thread> threads;
    const std::function<void(int)> *job = nullptr;
    int next = 0, total = 0, pending = 0;
    uint64_t generation = 0;
};

// host-to-host copy of a (rows x cols) block, split over the copy pool when it is large
void host_copy_rows(double *dst, size_t dst_ld, const double *src, size_t src_ld, size_t cols,
                    size_t rows)
{
    auto part = [=](size_t r0, size_t r1) {
        if (dst_ld == cols && src_ld == cols) {
            memcpy(dst + r0 * cols, src + r0 * cols, (r1 - r0) * cols * 8);
        } else {
            for (size_t r = r0; r < r1; ++r) memcpy(dst + r * dst_ld, src + r * src_ld, cols * 8);
        }
    };
    const size_t bytes = rows * cols * 8;
    CopyPool &pool = CopyPool::get();
    const size_t nt = std::min<size_t>({(size_t)pool.size(), bytes / ((size_t)1 << 20), rows});
    if (nt <= 1) {
        part(0, rows);
        return;
    }
    // twice as many pieces as threads: the tail is short when a thread starts late
    const int parts = (int)std::min<size_t>(2 * nt, rows);
    pool.run(parts, [&](int p) { part(rows * p / parts, rows * (p + 1) / parts); });
}

int ensure_host_streams(mpx_plan *pl)
{
    for (int b = 0; b < 2; ++b) {
        if (!pl->h_stream[b]) MPX_CUDA(cudaStreamCreateWithFlags(&pl->h_stream[b], cudaStreamNonBlocking));
        if (!pl->h_done[b]) MPX_CUDA(cudaEventCreateWithFlags(&pl->h_done[b], cudaEventDisableTiming));
        if (!pl->h_up[b]) MPX_CUDA(cudaEventCreateWithFlags(&pl->h_up[b], cudaEventDisableTiming));
    }
    if (!pl->h_event) MPX_CUDA(cudaEventCreateWithFlags(&pl->h_event, cudaEventDisableTiming));
    return MPX_OK;
}

#define MPX_TRY(expr)                                                           \
    do {                                                                        \
        cudaError_t _e = (expr);                                                \
        if (_e != cudaSuccess) {                                                \
            status = mpx_fail(MPX_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
            goto done;                                                          \
        }                                                                       \
    } while (0)

}  // namespace

// Evaluation with host buffers.  The coefficients are uploaded and packed once (into an image of
// their own: a tensor-path call running on the same plan keeps its coefficients); the points go
// through two device blocks on two streams, so that the upload of block b+1 and the download of
// block b-1 overlap the evaluation of block b.  Pageable caller memory is staged through
// page-locked blocks by host threads, one block ahead of (points) / behind (values) the device.
// Staging buffers and streams belong to the plan and are reused by later calls.
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
    std::lock_guard<std::mutex> lock(pl->host_mu);

    // block size: ~64 MiB of points+values per block, at least 64 Ki points
    int64_t block = ((int64_t)64 << 20) / (8 * (m + k));
    block = std::max<int64_t>(block, (int64_t)1 << 16);
    block = std::min(block, P);
    const int nbuf = (P > block) ? 2 : 1;
    // (below 16 MiB the driver's own staging of a pageable copy is as fast as ours)
    const bool stage_x = is_pageable(xx_h) && (size_t)P * m * 8 >= ((size_t)16 << 20);
    const bool stage_o = is_pageable(out_h) && (size_t)P * k * 8 >= ((size_t)16 << 20);
    if ((rc = ensure_host_streams(pl)) != MPX_OK) return rc;
    if ((rc = grow(&pl->h_c, &pl->h_c_cap, (size_t)pl->N * k * 8)) != MPX_OK) return rc;
    for (int b = 0; b < nbuf; ++b) {
        if ((rc = grow(&pl->h_x[b], &pl->h_x_cap[b], (size_t)block * m * 8)) != MPX_OK) return rc;
        if ((rc = grow(&pl->h_o[b], &pl->h_o_cap[b], (size_t)block * k * 8)) != MPX_OK) return rc;
        if (stage_x && (rc = grow_pinned(&pl->p_x[b], &pl->p_x_cap[b], (size_t)block * m * 8)) != MPX_OK) return rc;
        if (stage_o && (rc = grow_pinned(&pl->p_o[b], &pl->p_o_cap[b], (size_t)block * k * 8)) != MPX_OK) return rc;
    }

    cudaStream_t s0 = pl->h_stream[0];
    MPX_CUDA(copy_rows(pl->h_c, k, coeffs_h, ld_c, k, pl->N, cudaMemcpyHostToDevice, s0));
    if ((rc = mpx_launch_eval(pl, &pl->host_image, nullptr, 0, m, pl->h_c, k, k, nullptr, k, s0)) != MPX_OK) return rc;
    MPX_CUDA(cudaEventRecord(pl->h_event, s0));
    MPX_CUDA(cudaStreamWaitEvent(pl->h_stream[1], pl->h_event, 0));

    int status = MPX_OK;
    const int64_t nblk = (P + block - 1) / block;
    auto drain = [&](int64_t b) {          // staged values of block b: wait for its D2H, copy out
        const int s = (int)(b % nbuf);
        const int64_t p0 = b * block, np = std::min(block, P - p0);
        cudaError_t e = cudaEventSynchronize(pl->h_done[s]);
        if (e != cudaSuccess) return e;
        host_copy_rows(out_h + p0 * ld_out, ld_out, pl->p_o[s], k, k, np);
        return cudaSuccess;
    };
    for (int64_t b = 0; b < nblk; ++b) {
        const int s = (int)(b % nbuf);
        cudaStream_t st = pl->h_stream[s];
        const int64_t p0 = b * block, np = std::min(block, P - p0);
        if (stage_x) {
            // the page-locked block is free once the upload of block b - nbuf has finished
            if (b >= nbuf) MPX_TRY(cudaEventSynchronize(pl->h_up[s]));
            host_copy_rows(pl->p_x[s], m, xx_h + p0 * ld_xx, ld_xx, m, np);
            MPX_TRY(cudaMemcpyAsync(pl->h_x[s], pl->p_x[s], (size_t)np * m * 8, cudaMemcpyHostToDevice, st));
            MPX_TRY(cudaEventRecord(pl->h_up[s], st));
        } else {
            MPX_TRY(copy_rows(pl->h_x[s], m, xx_h + p0 * ld_xx, ld_xx, m, np, cudaMemcpyHostToDevice, st));
        }
        status = mpx_launch_eval(pl, &pl->host_image, pl->h_x[s], np, m, nullptr, k, k, pl->h_o[s], k, st);
        if (status != MPX_OK) goto done;
        if (stage_o) {
            MPX_TRY(cudaMemcpyAsync(pl->p_o[s], pl->h_o[s], (size_t)np * k * 8, cudaMemcpyDeviceToHost, st));
            MPX_TRY(cudaEventRecord(pl->h_done[s], st));
            if (b >= 1) MPX_TRY(drain(b - 1));       // before block b + 1 reuses that buffer
        } else {
            MPX_TRY(copy_rows(out_h + p0 * ld_out, ld_out, pl->h_o[s], k, k, np, cudaMemcpyDeviceToHost, st));
        }
    }
    if (stage_o) MPX_TRY(drain(nblk - 1));
done:
    for (int b = 0; b < 2; ++b) {
        cudaError_t e = cudaStreamSynchronize(pl->h_stream[b]);
        if (e != cudaSuccess && status == MPX_OK) status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e));
    }
    return status;
}

// Divided differences with host buffers.  The value array lives in a device buffer owned by the
// plan (no allocation per call).  The kernels take 0.1-0.3 ms of a call that moves 2 x 122 MB at
// cfg4, so the call is shaped by the transfers.
//   * page-locked arrays: columns are independent, so the array travels in blocks of columns on
//     two streams (2-D copies run at the full 55 GB/s, tools/copy2d_probe.cu): the upload of block
//     j+1, the kernels of block j and the download of block j-1 overlap, the bus being full duplex
//     -- 3.2 ms for 2 x 122 MB;
//   * a pageable array is repacked through page-locked staging by the copy pool.  That keeps up
//     with the bus only for contiguous pieces (55 GB/s; the 512-byte column pieces of the scheme
//     above reached 19 GB/s in place; page-locking the array with cudaHostRegister takes 16 ms per
//     122 MB -- tools/host_copy_probe.cpp, tools/hostreg_probe.cu), so these calls move blocks of
//     whole rows: upload pipelined with the repacking, kernels over the whole array, download
//     pipelined with the copy-out.
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
    std::lock_guard<std::mutex> lock(pl->host_mu);
    const int64_t N = pl->N;
    if ((rc = ensure_host_streams(pl)) != MPX_OK) return rc;
    if ((rc = grow(&pl->h_v, &pl->h_v_cap, (size_t)N * k * 8)) != MPX_OK) return rc;

    const bool big = (size_t)N * k * 8 >= ((size_t)16 << 20);
    const bool stage_i = big && is_pageable(in_h);
    const bool stage_o = big && is_pageable(out_h);
    int status = MPX_OK;
    if (!stage_i && !stage_o) {
        // page-locked (or small) arrays: blocks of columns, multiples of 64 (what one warp of the
        // kernel takes), about a quarter of the array each; a single block when it is small
        int kb = k;
        if ((size_t)N * k * 8 >= ((size_t)8 << 20) && k > 64) kb = std::max(64, ((k + 3) / 4 + 63) / 64 * 64);
        const int nblk = (k + kb - 1) / kb;
        for (int j = 0; j < nblk; ++j) {
            const int c0 = j * kb, kc = std::min(kb, k - c0);
            cudaStream_t st = pl->h_stream[j % 2];
            double *d_blk = pl->h_v + c0;                     // columns c0.. of the (N x k) array
            MPX_TRY(copy_rows(d_blk, k, in_h + c0, ld_in, kc, N, cudaMemcpyHostToDevice, st));
            status = mpx_launch_dds(pl, d_blk, kc, k, st);
            if (status != MPX_OK) goto done;
            MPX_TRY(copy_rows(out_h + c0, ld_out, d_blk, k, kc, N, cudaMemcpyDeviceToHost, st));
        }
    } else {
        // a pageable array on either side: blocks of whole rows, about 32 MiB, through
        // page-locked staging (two blocks in flight)
        const int64_t rb = std::max<int64_t>(1, std::min<int64_t>(N, ((int64_t)32 << 20) / ((int64_t)k * 8)));
        const int64_t nblk = (N + rb - 1) / rb;
        const int nbuf = nblk > 1 ? 2 : 1;
        for (int b = 0; b < nbuf; ++b) {
            if (stage_i && (rc = grow_pinned(&pl->p_x[b], &pl->p_x_cap[b], (size_t)rb * k * 8)) != MPX_OK) return rc;
            if (stage_o && (rc = grow_pinned(&pl->p_o[b], &pl->p_o_cap[b], (size_t)rb * k * 8)) != MPX_OK) return rc;
        }
        cudaStream_t st = pl->h_stream[0];
        auto drain = [&](int64_t b) {
            const int s = (int)(b % nbuf);
            const int64_t r0 = b * rb, nr = std::min(rb, N - r0);
            cudaError_t e = cudaEventSynchronize(pl->h_done[s]);
            if (e != cudaSuccess) return e;
            host_copy_rows(out_h + r0 * ld_out, ld_out, pl->p_o[s], k, k, nr);
            return cudaSuccess;
        };
        if (stage_i) {
            for (int64_t b = 0; b < nblk; ++b) {
                const int s = (int)(b % nbuf);
                const int64_t r0 = b * rb, nr = std::min(rb, N - r0);
                if (b >= nbuf) MPX_TRY(cudaEventSynchronize(pl->h_up[s]));
                host_copy_rows(pl->p_x[s], k, in_h + r0 * ld_in, ld_in, k, nr);
                MPX_TRY(cudaMemcpyAsync(pl->h_v + r0 * k, pl->p_x[s], (size_t)nr * k * 8, cudaMemcpyHostToDevice, st));
                MPX_TRY(cudaEventRecord(pl->h_up[s], st));
            }
        } else {
            MPX_TRY(copy_rows(pl->h_v, k, in_h, ld_in, k, N, cudaMemcpyHostToDevice, st));
        }
        status = mpx_launch_dds(pl, pl->h_v, k, k, st);
        if (status != MPX_OK) goto done;
        if (stage_o) {
            for (int64_t b = 0; b < nblk; ++b) {
                const int s = (int)(b % nbuf);
                const int64_t r0 = b * rb, nr = std::min(rb, N - r0);
                MPX_TRY(cudaMemcpyAsync(pl->p_o[s], pl->h_v + r0 * k, (size_t)nr * k * 8, cudaMemcpyDeviceToHost, st));
                MPX_TRY(cudaEventRecord(pl->h_done[s], st));
                if (b >= 1) MPX_TRY(drain(b - 1));
            }
            MPX_TRY(drain(nblk - 1));
        } else {
            MPX_TRY(copy_rows(out_h, ld_out, pl->h_v, k, k, N, cudaMemcpyDeviceToHost, st));
        }
    }
done:
    for (int b = 0; b < 2; ++b) {
        cudaError_t e = cudaStreamSynchronize(pl->h_stream[b]);
        if (e != cudaSuccess && status == MPX_OK) status = mpx_fail(MPX_ERR_CUDA, cudaGetErrorString(e));
    }
    return status;
}
