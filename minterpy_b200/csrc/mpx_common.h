// Internal declarations shared by the translation units of libminterpy_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/minterpy_b200.h"

// ---- layout of the evaluation "program" (see DESIGN.md, section "Evaluation kernel") ----
// The index set is compiled once per plan into chunks.  One chunk = a fixed-size block of
// leaf records (descriptor words) plus the coefficient slots those leaves consume, so a CTA
// can stream chunk after chunk through shared memory with two bulk copies per chunk.
constexpr int MPX_DESC_WORDS = 512;   // u32 per chunk: [0]=number of words used, [4..] records
constexpr int MPX_DESC_HDR = 4;
constexpr int MPX_COEF_CAP = 1024;    // coefficient slots (doubles) per chunk and polynomial
constexpr int MPX_STAGES = 8;         // shared-memory ring depth of the chunk stream
constexpr int MPX_MAX_L0 = 66;        // dimension-0 table entries a thread can hold in registers
constexpr int MPX_MAX_M = 16;
constexpr int MPX_DDS_REG_LEN = 16;   // longest fiber the register-resident dds kernel takes
constexpr int MPX_DDS_COLS = 32;      // value columns per CTA of the fused dds kernel
constexpr int MPX_DDS_MERGE_ROWS = 112; // small slices are merged until a CTA owns this many rows
constexpr int MPX_DDS_MAX_ROWS = 736; // rows of a slice that fit shared memory (x 32 columns x 8 B)

// Record = one header word, MPX count words, then extension words when MPX_D_IRREG is set.
// A record covers consecutive leaves (dimension-1 exponent counting upwards from e1_first)
// whose padded lengths do not increase; the count words say how many of them have K
// coefficient pairs, 4 bits per K (K = 1 in the lowest nibble), at most 15 each.  The kernel
// walks K downwards through straight-line code specialised for each K.
//   header bits 0..7   dimension-1 exponent of the first leaf (table row of T1)
//          bits 8..11  nf: how many upper levels (dimensions 2..) receive a fold afterwards
//          bit  12     irregular: advance/restart amounts follow in extension words
//          bit  13     last record of the program (no advance after the folds)
//          bits 14..21 exponent of the advancing level before it advances (regular case)
constexpr uint32_t MPX_D_IRREG = 1u << 12;
constexpr uint32_t MPX_D_LAST = 1u << 13;

// kernel variant = size of the dimension-0 register table (even, >= max exponent + 1)
inline int mpx_l0p_for(int max_exp0)
{
    const int L0 = max_exp0 + 1;
    return L0 <= 10 ? 10 : L0 <= 18 ? 18 : L0 <= 26 ? 26 : L0 <= 34 ? 34 : 66;
}
inline int mpx_count_words(int l0p) { return (l0p / 2 + 7) / 8; }

// The used words of the record stream, for kernels that take the program as a KERNEL PARAMETER
// (constant bank, read through the uniform datapath): per chunk the header and the words it
// uses, chunk c at w[off[c]].  16.9 KB of the 32 KB a launch may carry.
constexpr int MPX_U_WORDS = 4096;
constexpr int MPX_U_CHUNKS = 127;
struct MpxUniformProgram {
    uint32_t off[MPX_U_CHUNKS + 1];
    uint32_t w[MPX_U_WORDS];
};

struct EvalProgram {
    int n_chunks = 0;
    int64_t n_leaves = 0;
    int64_t fma_per_point = 0;     // leaf FMAs (padded) + level-1 FMAs + folds
    int64_t other_fp64_per_point = 0;  // DADD/DMUL of running products and the tables
    std::vector<uint32_t> desc;    // n_chunks * MPX_DESC_WORDS
    std::vector<int32_t> slot_src; // n_chunks * MPX_COEF_CAP, row of the coefficient or -1
};

// ---- tile program of the tensor-core tree kernel (mpx_eval_tmma.cu) ------------------------
// For complete index sets the leaf sums  sum_e0 c[leaf][e0] * T0[point][e0]  form a dense
// contraction [points x e0] * [e0 x leaves] that DMMA m8n8k4 computes 8 leaves at a time.  A
// TILE is 8 columns (leaves with consecutive dimension-1 exponents; the columns may continue
// into the head of the next (e2.., e_{m-1}) group) and `ks` k-steps of 4 dimension-0
// exponents; its coefficients are stored as B fragments, two k-steps interleaved so that one 16-byte
// load feeds both: slot = (s/2)*64 + (column*4 + (e0 & 3))*2 + (s & 1).
// Tile record, two words:
//   w0 bits 0..4 ks | 5..8 columns of the current group | 9..12 columns of the next group |
//      13..20 dimension-1 exponent of column 0 | 21 the current group ends in this tile |
//      22..25 nf (folds into dimensions 2.. when it ends) | 26 last tile of the program
//   w1 exponent (before advancing) of the level that advances after the fold
constexpr int MPX_T_DESC_WORDS = 128;   // u32 per chunk: [0] = words used, [4..] records
constexpr int MPX_T_COEF_CAP = 2048;    // coefficient slots per chunk (64 k-steps)
struct TileProgram {
    bool usable = false;          // built (complete set, m >= 2) and not wasteful
    int n_chunks = 0;
    int ks_max = 0;               // k-steps the dimension-0 table needs: ceil((max_exp0+1)/4)
    int64_t n_tiles = 0;
    int64_t mma_fma_per_point = 0;    // 32 * sum of ks over the tiles (tensor FMAs per point)
    int64_t fma_per_point = 0;        // + level-1 FMAs (8 per tile) + folds
    std::vector<uint32_t> desc;       // n_chunks * MPX_T_DESC_WORDS
    std::vector<int32_t> slot_src;    // n_chunks * MPX_T_COEF_CAP
};

struct DdsDim {
    int n_fibers = 0;
    int max_len = 0;
    int32_t *d_off = nullptr;   // n_fibers + 1
    int32_t *d_rows = nullptr;  // N, rows grouped by fiber, in order of increasing exponent
    double *d_diff = nullptr;   // (L x L) table: diff[l*L + r] = gp[r,d] - gp[l,d], L = max_len
    // work list of the one-warp-per-fiber kernel: fibers with at least 2 elements, longest first
    int n_work = 0;
    int32_t *d_woff = nullptr;  // n_work + 1
    int32_t *d_wrows = nullptr; // rows of those fibers
    int32_t *d_wpad = nullptr;  // the same rows, max_len slots per fiber (column kernel: the row
                                // numbers can be read before the fiber's length is known)
    double *d_dy = nullptr;     // (L x L) pairs (divisor, reciprocal or 0), 2 doubles each
    double *d_gcol = nullptr;   // (L) generating points of this dimension
    // several-columns-per-lane kernel (max_len <= 40): (divisor, reciprocal) pairs as d_dy, except
    // that a divisor whose significand is all ones -- 2^j (1 - 2^-53) up to sign, where the
    // reciprocal-and-correct division is not exact -- is replaced by the pair (+-2^j, +-2^-j) and
    // flagged in d_fmask (bit r of word l): its quotient is the next double after a / 2^j, away
    // from zero (see mpx_dds.cu).  cols_ok: every divisor is ordinary or of that kind.
    double *d_dy2 = nullptr;
    unsigned long long *d_fmask = nullptr;
    bool cols_ok = false;
    // register-fiber kernel (max_len <= MPX_DDS_REG_LEN): the pairs (divisor, reciprocal or 0)
    // for 0 <= l < r < max_len at [r (r - 1) / 2 + l], handed to the kernel as a parameter so
    // that every table entry is an immediate constant-bank operand of the FP64 instructions
    std::vector<double> h_pairs;
    bool all_fast = false;      // every divisor of this dimension qualifies for the fast division
};

// A group of consecutive dimensions [lo, hi] that the divided-difference kernel handles in
// one pass: the index set falls apart into SLICES (rows that agree in every dimension outside
// the group); a slice is loaded into shared memory, all fibers of dimensions hi..lo inside it
// are processed there, and it is written back.
struct DdsGroupDim {
    int dim = 0;
    int L = 0;                      // longest fiber (side of the diff table)
    int max_fibers = 0;             // most fibers of this dimension in one slice
    int32_t *d_slice_fib0 = nullptr;  // n_slices + 1: first fiber of every slice
    int32_t *d_fib_off = nullptr;     // n_fibers + 1: offsets into d_fib_rows
    uint16_t *d_fib_rows = nullptr;   // N: slice-local row of every fiber element, e_d ascending
    double *d_diff = nullptr;         // (L x L): diff[l*L + r] = gp[r,dim] - gp[l,dim]
    double *d_recip = nullptr;        // (L x L): 1.0 / diff, correctly rounded (host division)
    double *d_g = nullptr;            // (L): generating points of this dimension
};
struct DdsGroup {
    int lo = 0, hi = 0;
    int n_slices = 0;
    int max_rows = 0;               // largest slice
    int32_t *d_slice_off = nullptr; // n_slices + 1
    int32_t *d_slice_rows = nullptr;// N: global row of every slice element
    std::vector<DdsGroupDim> dims;  // hi first
};

// One packed-coefficient image in device memory and what orders its users on the device.
// The evaluation kernels read the image long after the entry point has returned, so before the
// next call repacks it on ANOTHER stream it waits for `busy`, the event recorded behind the last
// kernel that read it (same stream: stream order already does it).
struct MpxImage {
    double *ptr = nullptr;
    size_t bytes = 0;
    bool fixed = false;              // caller-provided storage of fixed size: never grown or freed
    cudaEvent_t busy = nullptr;      // created on first use (not for fixed images)
    cudaStream_t last_stream = nullptr;
    bool in_use = false;             // `busy` has been recorded at least once
};

struct mpx_plan {
    int device = 0;
    int64_t N = 0;
    int m = 0;
    int gp_rows = 0;                 // rows kept (max exponent over all dims + 1)
    int max_exp[MPX_MAX_M] = {0};
    int start_exp[MPX_MAX_M] = {0};  // exponents of the first leaf in dimensions >= 2
    bool complete = false;
    bool tree_ok = true;             // dimension-0 exponents fit the register table of the tree kernel
    // evaluation
    EvalProgram prog;
    std::vector<MpxUniformProgram> uprog;   // one element if the record stream qualifies, else empty
    TileProgram tprog;               // tensor-core tree kernel (complete sets)
    uint32_t *d_tdesc = nullptr;
    int32_t *d_tslot_src = nullptr;
    uint32_t *d_desc = nullptr;
    int32_t *d_slot_src = nullptr;
    int32_t *d_row1d = nullptr;      // m = 1, distinct exponents: row of the monomial with exponent e, or -1
    double *d_gp = nullptr;          // (m x gp_rows) dimension-major copy of the generating points
    uint8_t *d_exp8 = nullptr;       // (N x m) exponents as bytes, for the monomial kernel
    int *d_max_exp = nullptr;        // (m) largest exponent per dimension
    uint64_t *d_exp64 = nullptr;     // exponents packed 8 bytes per monomial, padded to 32 rows (m <= 8)
    int mma_chunks = 0;              // ceil(N / 32)
    // dds
    std::vector<DdsDim> dds;         // one per dimension (only if complete)
    std::vector<DdsGroup> dds_groups; // fused passes, top dimensions first (only if complete)
    bool dds_fast_div = true;         // every divisor qualifies for the reciprocal + correction division
    int64_t dds_updates = 0;
    // packed coefficients of the evaluation call in flight, one image per way in: the
    // device-pointer entry point (mpx_eval, mpx_eval_deriv) and the host-buffer entry point each
    // own one, so a NumPy call can never repack what a tensor call is still reading
    std::mutex mu;                   // host-side state of `image` (pointer, size, event)
    MpxImage image;                  // mpx_eval / mpx_eval_deriv
    MpxImage deriv_tables;           // mpx_eval_deriv on the tree kernel: tables of dimensions >= 2
    std::mutex host_mu;              // one host-buffer call per plan at a time (held to the end)
    MpxImage host_image;             // mpx_eval_host
    // staging for the host-buffer entry points: two streams, device blocks for points and
    // values, page-locked blocks for callers whose arrays are pageable, the dds value array
    cudaStream_t h_stream[2] = {nullptr, nullptr};
    cudaEvent_t h_event = nullptr;
    cudaEvent_t h_done[2] = {nullptr, nullptr};       // block b's D2H has finished
    cudaEvent_t h_up[2] = {nullptr, nullptr};         // block b's H2D has finished
    double *h_x[2] = {nullptr, nullptr}, *h_o[2] = {nullptr, nullptr}, *h_c = nullptr;
    size_t h_x_cap[2] = {0, 0}, h_o_cap[2] = {0, 0}, h_c_cap = 0;   // bytes
    double *p_x[2] = {nullptr, nullptr}, *p_o[2] = {nullptr, nullptr};   // page-locked host blocks
    size_t p_x_cap[2] = {0, 0}, p_o_cap[2] = {0, 0};
    double *h_v = nullptr;           // dds_host: the value array on the device
    size_t h_v_cap = 0;
};

// error plumbing --------------------------------------------------------------------------
int mpx_fail(int code, const std::string &msg);
#define MPX_CUDA(expr)                                                                     \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess)                                                             \
            return mpx_fail(MPX_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)

// Dynamic shared memory limit of a kernel: always raised to the device's opt-in maximum, never
// to the size of one launch.  The attribute belongs to the function, process-wide, so two host
// threads launching the same kernel with different sizes (two plans) would otherwise race: one
// lowers the limit between the other's set and launch.
// Done once per (kernel, device): the three runtime calls cost ~2 us of every launch otherwise.
bool mpx_smem_raised(const void *fn, int device, bool mark);       // mpx_plan.cu
template <typename F>
inline cudaError_t mpx_allow_max_smem(F fn)
{
    int device = 0, smem_max = 0;
    cudaError_t e = cudaGetDevice(&device);            // the entry points select the plan's device
    if (e != cudaSuccess) return e;
    if (mpx_smem_raised(reinterpret_cast<const void *>(fn), device, false)) return cudaSuccess;
    e = cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
    if (e == cudaSuccess) mpx_smem_raised(reinterpret_cast<const void *>(fn), device, true);
    return e;
}

struct DeviceGuard {
    int prev = -1;
    bool ok = false, changed = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) return;
        if (prev == dev) { ok = true; return; }          // the usual case: nothing to switch
        if (cudaSetDevice(dev) == cudaSuccess) ok = changed = true;
    }
    ~DeviceGuard() { if (changed) cudaSetDevice(prev); }
};

// kernels' host launchers (defined in the .cu files) ---------------------------------------
// All evaluation launchers: `img` receives the packed coefficients (coeffs != nullptr) and is
// what the kernel reads; coeffs == nullptr means `img` already holds these k polynomials.
int mpx_launch_eval(mpx_plan *plan, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                    const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                    cudaStream_t stream);
int mpx_launch_monomials(mpx_plan *plan, const double *xx, int64_t P, int64_t ld_xx,
                         double *out, int64_t ld_out, int triangular, cudaStream_t stream);
int mpx_launch_dds(mpx_plan *plan, double *vals, int k, int64_t ld, cudaStream_t stream, int flags = 0);
// make room for `bytes` in the image and order `stream` behind the image's last reader
int mpx_image_acquire(MpxImage *img, size_t bytes, cudaStream_t stream);
// record that work reading the image has just been enqueued on `stream`
int mpx_image_release(MpxImage *img, cudaStream_t stream);
void mpx_image_destroy(MpxImage *img);
// tensor-core (many polynomials) evaluation path, mpx_eval_mma.cu
bool mpx_mma_eligible(const mpx_plan *plan, int k);
int mpx_mma_min_k(const mpx_plan *plan);      // smallest k that prefers this kernel to the tree kernel
size_t mpx_mma_scratch_bytes(const mpx_plan *plan, int k);
// orders_h (m entries, host) != nullptr: evaluate the partial derivative of those orders
int mpx_launch_eval_mma(mpx_plan *plan, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                        const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                        cudaStream_t stream, const int *orders_h = nullptr);
bool mpx_mma_fits(const mpx_plan *plan, int k);   // the tables of this index set fit shared memory
// tensor-core tree kernel (complete sets, few polynomials), mpx_eval_tmma.cu
// one-dimensional polynomials by Horner's rule (mpx_eval_1d.cu)
bool mpx_1d_eligible(const mpx_plan *plan, int k);
size_t mpx_1d_scratch_bytes(const mpx_plan *plan, int k);
int mpx_launch_eval_1d(mpx_plan *plan, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                       const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                       cudaStream_t stream);
bool mpx_tmma_eligible(const mpx_plan *plan, int k);
size_t mpx_tmma_scratch_bytes(const mpx_plan *plan, int k);
int mpx_launch_eval_tmma(mpx_plan *plan, MpxImage *img, const double *xx, int64_t P, int64_t ld_xx,
                         const double *coeffs, int k, int64_t ld_c, double *out, int64_t ld_out,
                         cudaStream_t stream);
// few polynomials: partial derivative on the tree kernel, mpx_eval_deriv.cu
int mpx_launch_eval_tree_deriv(mpx_plan *plan, MpxImage *img, MpxImage *tables, const double *xx,
                               int64_t P, int64_t ld_xx, const double *coeffs, int k, int64_t ld_c,
                               const int *orders_h, double *out, int64_t ld_out, cudaStream_t stream);
extern int mpx_test_eval_stages;   // test hook, mpx_eval.cu
extern int mpx_test_eval_uniform;  // test hook, mpx_eval.cu: 2 = never take the uniform-program kernels
// bytes of plan scratch an evaluation of k polynomials needs (either path)
size_t mpx_eval_scratch_bytes(const mpx_plan *plan, int k);
