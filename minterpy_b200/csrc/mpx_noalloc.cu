// mpx_eval_noalloc: polynomial evaluation for callers that must not allocate device memory.
//
// Who needs this: a host application that handed all device memory to its own pool when it
// started (a later cudaMalloc would simply fail), or one that evaluates inside a
// latency-critical loop and cannot accept the device-wide synchronisation that cudaMalloc and
// cudaFree imply.  mpx_eval may grow the plan's scratch buffer (the packed coefficient image)
// with cudaMalloc; this variant keeps that image in a fixed array that is part of the
// library's device image instead, so the call itself allocates nothing.
//
// Cost and limits: every GPU that loads libminterpy_b200.so holds MPX_NOALLOC_BYTES (8 MiB)
// for the array until the process exits.  A call with few polynomials (k below the plan's
// MPX_INFO_MMA_MIN_K) needs chunks * 1024 * k doubles (chunks = MPX_INFO_CHUNKS of the plan), one
// with more about 1.06 * N * k doubles; it
// is refused with MPX_ERR_UNSUPPORTED if that exceeds the array.  All plans and all calls on a device share the one array: it still holds the previous
// call's image when the next call starts (every slot the next call reads is rewritten first,
// padding included, so nothing has to be cleared), and two calls whose kernels overlap on the
// GPU -- different streams, different host threads -- would overwrite each other's
// coefficients.  Issue them so that one finishes before the next begins.
#include <mutex>

#include "mpx_common.h"

namespace {

constexpr size_t MPX_NOALLOC_SLOTS = (size_t)1 << 20;               // doubles
constexpr size_t MPX_NOALLOC_BYTES = MPX_NOALLOC_SLOTS * sizeof(double);

// The scratch itself.  In host code this name is only a handle for the runtime, not a device
// address: the address a kernel can use comes from cudaGetSymbolAddress (below).
static __device__ __align__(256) double g_noalloc_image[MPX_NOALLOC_SLOTS];

constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mutex;
double *g_image_ptr[MAX_DEVICES] = {nullptr};

// Once per device, with that device current: ask the runtime where the array lives there.
int noalloc_image(int device, double **ptr_out)
{
    if (device < 0 || device >= MAX_DEVICES) return mpx_fail(MPX_ERR_ARG, "device index out of range");
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_image_ptr[device]) {
        double *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_image);
        if (e == cudaErrorMemoryAllocation) {
            // with lazy module loading (the runtime's default) the device image that holds the
            // array is only placed on the GPU now; if the application has already taken all
            // device memory for its own pool there is no room left for it
            cudaGetLastError();
            return mpx_fail(MPX_ERR_NOMEM,
                            "mpx_eval_noalloc set-up: no device memory left for the library's 8 MiB "
                            "scratch array; make the first call on each device before the application "
                            "takes its memory pool, or run with CUDA_MODULE_LOADING=EAGER");
        }
        if (e != cudaSuccess || p == nullptr)
            return mpx_fail(MPX_ERR_CUDA, std::string("cudaGetSymbolAddress(no-alloc scratch): ") +
                                              cudaGetErrorString(e));
        g_image_ptr[device] = p;
    }
    *ptr_out = g_image_ptr[device];
    return MPX_OK;
}

}  // namespace

extern "C" int mpx_eval_noalloc(mpx_plan *pl, const double *xx, int64_t P, int64_t ld_xx,
                                const double *coeffs, int k, int64_t ld_c, double *out,
                                int64_t ld_out, void *stream)
{
    if (!pl) return mpx_fail(MPX_ERR_ARG, "plan is null");
    if (P < 0 || k < 1) return mpx_fail(MPX_ERR_ARG, "P must be >= 0 and k >= 1");
    if (P == 0) return MPX_OK;
    if (!xx || !coeffs || !out) return mpx_fail(MPX_ERR_ARG, "null device pointer");
    if (ld_xx < pl->m || ld_c < k || ld_out < k) return mpx_fail(MPX_ERR_ARG, "leading dimension too small");
    const size_t need = mpx_eval_scratch_bytes(pl, k);     // bytes
    if (need > MPX_NOALLOC_BYTES)
        return mpx_fail(MPX_ERR_UNSUPPORTED,
                        "mpx_eval_noalloc: the packed coefficients of this call need " +
                            std::to_string(need) + " bytes, the fixed scratch holds " +
                            std::to_string(MPX_NOALLOC_BYTES) + "; use fewer polynomials per call or mpx_eval");
    DeviceGuard dg(pl->device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select the plan's CUDA device");
    double *image = nullptr;
    int rc = noalloc_image(pl->device, &image);     // first call on this device: set-up
    if (rc != MPX_OK) return rc;

    // The regular launch sequence, with the fixed array as the image that receives the packed
    // coefficients.  It is marked fixed: it is never grown or freed, and it carries no event
    // (creating one would be an allocating call) -- overlapping calls are the caller's to avoid.
    MpxImage fixed_image;
    fixed_image.ptr = image;
    fixed_image.bytes = MPX_NOALLOC_BYTES;
    fixed_image.fixed = true;
    return mpx_launch_eval(pl, &fixed_image, xx, P, ld_xx, coeffs, k, ld_c, out, ld_out,
                           (cudaStream_t)stream);
}
