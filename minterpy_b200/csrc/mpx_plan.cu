// Plan construction: compiles the exponent matrix into (a) the chunked leaf program the
// evaluation kernel streams, (b) per-dimension fiber tables for the divided-difference
// kernel, and uploads them together with the generating points.  Host-side integer work;
// replaces the per-call preparation of src/minterpy/utils.py:218-231 and the tree metadata
// of src/minterpy/core/tree.py:47-104 / src/minterpy/dds.py:30-235,473-602.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <numeric>
#include <set>
#include <utility>

#include "mpx_common.h"

static thread_local std::string g_last_error;

int mpx_fail(int code, const std::string &msg)
{
    g_last_error = msg;
    return code;
}

extern "C" const char *mpx_last_error(void) { return g_last_error.c_str(); }
extern "C" int mpx_version(void) { return 100; }

// Test hook (not in the public header), read when a plan is built: 0 = choose by padding
// overhead, 1 = tensor-core tree kernel whenever it can be built, 2 = never.
static int mpx_test_tmma = 0;
extern "C" void mpx_debug_set_tmma(int mode) { mpx_test_tmma = mode; }

namespace {

// lexicographic order with the LAST column most significant (np.lexsort(exponents.T),
// src/minterpy/core/utils.py:44-92)
struct LexLess {
    const int64_t *e;
    int m;
    bool operator()(int64_t a, int64_t b) const
    {
        for (int i = m - 1; i >= 0; --i) {
            int64_t x = e[a * m + i], y = e[b * m + i];
            if (x != y) return x < y;
        }
        return false;
    }
};

// The reciprocal + one correction division (mpx_dds.cu) returns the IEEE quotient only for
// ordinary divisors: moderate exponent, significand not all ones (e.g. 1 - 6e-17 between
// Chebyshev points DOES occur).  Others get reciprocal 0, which sends the kernel to the true
// division.
double reciprocal_or_zero(double dv)
{
    uint64_t bits;
    std::memcpy(&bits, &dv, 8);
    const int ex = (int)((bits >> 52) & 0x7ff);
    const uint64_t mant = bits & 0xfffffffffffffull;
    // |d| in [2^-100, 2^100]: with numerators in [2^-900, 2^900) neither the quotient nor the
    // remainder a - d q can leave the normal range (a wider allowance could overflow a * y)
    const bool ordinary = ex >= 1023 - 100 && ex <= 1023 + 100 && mant != 0xfffffffffffffull;
    return ordinary ? 1.0 / dv : 0.0;      // host division: RN(1/d)
}

// d = +-2^j (1 - 2^-53) with a moderate exponent: significand all ones.  Returns +-2^j in *pw.
bool all_ones_divisor(double dv, double *pw)
{
    uint64_t bits;
    std::memcpy(&bits, &dv, 8);
    const int ex = (int)((bits >> 52) & 0x7ff);
    const uint64_t mant = bits & 0xfffffffffffffull;
    if (mant != 0xfffffffffffffull || ex < 1023 - 100 || ex > 1023 + 100) return false;
    *pw = std::ldexp(dv < 0 ? -1.0 : 1.0, ex - 1023 + 1);
    return true;
}

int64_t find_sorted(const int64_t *e, const std::vector<int64_t> &order, int m,
                    const int64_t *key)
{
    int64_t lo = 0, hi = (int64_t)order.size() - 1;
    while (lo <= hi) {
        int64_t mid = (lo + hi) >> 1;
        const int64_t *row = e + order[mid] * m;
        int c = 0;
        for (int i = m - 1; i >= 0 && c == 0; --i) c = (row[i] < key[i]) ? -1 : (row[i] > key[i]);
        if (c == 0) return mid;
        if (c < 0) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

// ---------------------------------------------------------------------------------------
// Evaluation program.  Rows are visited in lexicographic order.  A LEAF is a maximal run
// of rows with equal exponents in dimensions 1.. and strictly increasing dimension-0
// exponent; its coefficients occupy slots 0..len-1 (len = largest e0 + 1, holes and the
// even-length padding get slot_src = -1 => coefficient 0).  Leaves are grouped into
// records (format: mpx_common.h); coefficient slots follow the order the kernel walks.
// ---------------------------------------------------------------------------------------
int compile_program(const int64_t *e, int64_t N, int m, const std::vector<int64_t> &order,
                    int l0p, EvalProgram &pg, std::string &err)
{
    const int U = std::max(m - 2, 0);
    const int ncw = mpx_count_words(l0p);
    const int cap_words = MPX_DESC_WORDS - MPX_DESC_HDR;
    struct Leaf { int64_t first, count; int len; };
    std::vector<Leaf> leaves;
    for (int64_t s = 0; s < N;) {
        int64_t t = s + 1;
        const int64_t *r0 = e + order[s] * m;
        while (t < N) {
            const int64_t *r1 = e + order[t] * m;
            bool same = true;
            for (int i = 1; i < m; ++i) same = same && (r1[i] == r0[i]);
            if (!same || r1[0] <= e[order[t - 1] * m]) break;
            ++t;
        }
        int len = (int)e[order[t - 1] * m] + 1;
        leaves.push_back({s, t - s, len});
        s = t;
    }
    pg.n_leaves = (int64_t)leaves.size();
    pg.fma_per_point = 0;
    pg.other_fp64_per_point = 0;
    pg.desc.clear();
    pg.slot_src.clear();
    pg.n_chunks = 0;
    pg.slot_src.resize(MPX_COEF_CAP, -1);

    std::vector<uint32_t> words;   // words of the current chunk
    int coef_used = 0;
    // the record under construction
    bool open = false;
    uint32_t hdr = 0;
    std::vector<uint32_t> counts(ncw, 0u), ext;
    int rec_e1_next = 0, rec_last_k = 0;

    auto close_record = [&]() {
        if (!open) return;
        words.push_back(hdr);
        for (uint32_t c : counts) words.push_back(c);
        for (uint32_t x : ext) words.push_back(x);
        open = false;
    };
    auto flush = [&]() {
        close_record();
        size_t base = pg.desc.size();
        pg.desc.resize(base + MPX_DESC_WORDS, 0u);
        pg.desc[base] = (uint32_t)words.size();
        std::copy(words.begin(), words.end(), pg.desc.begin() + base + MPX_DESC_HDR);
        pg.n_chunks++;
        pg.slot_src.resize((size_t)(pg.n_chunks + 1) * MPX_COEF_CAP, -1);   // room for the next chunk
        words.clear();
        coef_used = 0;
    };

    for (size_t li = 0; li < leaves.size(); ++li) {
        const Leaf &lf = leaves[li];
        const int64_t *cur = e + order[lf.first] * m;
        const int64_t *nxt = (li + 1 < leaves.size()) ? e + order[leaves[li + 1].first] * m : nullptr;
        const int lenp = (lf.len + 1) & ~1;
        const int K = lenp / 2;
        const int e1 = (m >= 2) ? (int)cur[1] : 0;
        // highest dimension in which the next leaf differs
        int D = 1;
        const bool last = (nxt == nullptr);
        if (last) {
            D = m - 1;
        } else {
            for (int i = m - 1; i >= 1; --i) if (nxt[i] != cur[i]) { D = i; break; }
        }
        const int nf = last ? U : std::max(D - 1, 0);
        uint32_t tail = (uint32_t)nf << 8;          // fold / advance fields of this leaf
        std::vector<uint32_t> leaf_ext;
        if (last) tail |= MPX_D_LAST;
        if (nf > 0 && !last) {
            // level u = nf-1 (dimension D) advances from cur[D] to nxt[D]; levels below restart
            int64_t adv = nxt[D] - cur[D];
            bool regular = (adv == 1);
            for (int i = 2; i < D; ++i) regular = regular && (nxt[i] == 0);
            if (regular) {
                tail |= (uint32_t)cur[D] << 14;
            } else {
                tail |= MPX_D_IRREG;
                leaf_ext.push_back((uint32_t)cur[D] | ((uint32_t)adv << 8));
                for (int i = 2; i < D; ++i) leaf_ext.push_back((uint32_t)nxt[i]);
                pg.other_fp64_per_point += 2 * (adv - 1);
                for (int i = 2; i < D; ++i) pg.other_fp64_per_point += 2 * nxt[i];
            }
            pg.other_fp64_per_point += 2;
        }
        // can the leaf join the open record?  next dimension-1 exponent, length not
        // increasing, nibble not full, and everything still fits into the chunk
        const bool fits_coef = coef_used + lenp <= MPX_COEF_CAP;
        bool join = open && fits_coef && e1 == rec_e1_next && K <= rec_last_k &&
                    ((counts[(K - 1) / 8] >> (4 * ((K - 1) % 8))) & 15u) < 15u &&
                    (int)(words.size() + 1 + ncw + leaf_ext.size()) <= cap_words;
        if (!join) {
            close_record();
            if (!fits_coef || (int)(words.size() + 1 + ncw + leaf_ext.size()) > cap_words) flush();
            open = true;
            hdr = (uint32_t)e1;
            std::fill(counts.begin(), counts.end(), 0u);
            ext.clear();
        }
        counts[(K - 1) / 8] += 1u << (4 * ((K - 1) % 8));
        rec_last_k = K;
        rec_e1_next = e1 + 1;
        // coefficient slots of this leaf
        int32_t *slots = pg.slot_src.data() + (size_t)pg.n_chunks * MPX_COEF_CAP + coef_used;
        for (int64_t r = 0; r < lf.count; ++r) {
            int64_t row = order[lf.first + r];
            slots[e[row * m]] = (int32_t)row;
        }
        coef_used += lenp;
        pg.fma_per_point += lenp + nf;      // lenp - 1 in the leaf sum, 1 into level 1
        // a leaf that folds or ends the program closes its record and hands it its tail
        if (nf > 0 || last) {
            hdr |= tail;
            ext = leaf_ext;
            close_record();
        }
    }
    flush();
    pg.slot_src.resize((size_t)pg.n_chunks * MPX_COEF_CAP);
    (void)err;
    return MPX_OK;
}

// ---------------------------------------------------------------------------------------
// Tile program for the tensor-core tree kernel (format: mpx_common.h).  Complete, sorted sets
// only: inside a group (rows that agree in dimensions 2..) the leaves have dimension-1
// exponents 0, 1, 2, .. and every leaf is a gap-free run e0 = 0..len-1.
// ---------------------------------------------------------------------------------------
void compile_tile_program(const int64_t *e, int64_t N, int m, TileProgram &tp)
{
    struct Leaf { int64_t first; int len; int group; };
    std::vector<Leaf> leaves;
    std::vector<int> group_end;    // index one past the last leaf of every group
    for (int64_t s = 0; s < N;) {
        int64_t t = s + 1;
        while (t < N) {
            bool same = true;
            for (int i = 1; i < m; ++i) same = same && (e[t * m + i] == e[s * m + i]);
            if (!same) break;
            ++t;
        }
        bool new_group = leaves.empty();
        if (!leaves.empty())
            for (int i = 2; i < m; ++i) new_group = new_group || (e[s * m + i] != e[leaves.back().first * m + i]);
        if (new_group && !leaves.empty()) group_end.push_back((int)leaves.size());
        leaves.push_back({s, (int)(t - s), (int)group_end.size()});
        s = t;
    }
    group_end.push_back((int)leaves.size());
    const int U = std::max(m - 2, 0);
    int max0 = 0;
    for (int64_t j = 0; j < N; ++j) max0 = std::max<int>(max0, (int)e[j * m]);
    tp.ks_max = (max0 + 1 + 3) / 4;
    tp.desc.clear();
    tp.slot_src.assign(MPX_T_COEF_CAP, -1);
    tp.n_chunks = 0;
    tp.n_tiles = 0;
    tp.mma_fma_per_point = tp.fma_per_point = 0;
    std::vector<uint32_t> words;
    int coef_used = 0;
    auto flush = [&]() {
        size_t base = tp.desc.size();
        tp.desc.resize(base + MPX_T_DESC_WORDS, 0u);
        tp.desc[base] = (uint32_t)words.size();
        std::copy(words.begin(), words.end(), tp.desc.begin() + base + MPX_DESC_HDR);
        tp.n_chunks++;
        tp.slot_src.resize((size_t)(tp.n_chunks + 1) * MPX_T_COEF_CAP, -1);
        words.clear();
        coef_used = 0;
    };
    const int L = (int)leaves.size();
    for (int i = 0; i < L;) {
        const int g = leaves[i].group;
        const int gend = group_end[g];
        const int ncur = std::min(8, gend - i);
        const bool ends = (i + ncur == gend);
        int nnext = 0;
        if (ends && ncur < 8 && gend < L) {
            // head of the next group, but never its last leaf (one group end per tile)
            const int next_size = group_end[g + 1] - gend;
            nnext = std::min(8 - ncur, next_size - 1);
        }
        int maxlen = 0;
        for (int c = 0; c < ncur + nnext; ++c) maxlen = std::max(maxlen, leaves[i + c].len);
        const int ks = (maxlen + 3) / 4;
        const bool last = ends && gend == L;
        const int64_t *cur = e + leaves[i].first * m;
        int nf = 0, adv_exp = 0;
        if (ends) {
            if (last) {
                nf = U;
            } else {
                const int64_t *nxt = e + leaves[gend].first * m;
                int D = 2;
                for (int d = m - 1; d >= 2; --d) if (nxt[d] != cur[d]) { D = d; break; }
                nf = D - 1;
                adv_exp = (int)cur[D];
            }
        }
        const int kpairs = (ks + 1) / 2;     // k-steps are stored in pairs: one LDS.128 feeds two
        if (coef_used + kpairs * 64 > MPX_T_COEF_CAP ||
            (int)words.size() + 2 > MPX_T_DESC_WORDS - MPX_DESC_HDR)
            flush();
        int32_t *slots = tp.slot_src.data() + (size_t)tp.n_chunks * MPX_T_COEF_CAP + coef_used;
        for (int c = 0; c < ncur + nnext; ++c) {
            const Leaf &lf = leaves[i + c];
            for (int e0 = 0; e0 < lf.len; ++e0)
                slots[(e0 >> 3) * 64 + (c * 4 + (e0 & 3)) * 2 + ((e0 >> 2) & 1)] = (int32_t)(lf.first + e0);
        }
        coef_used += kpairs * 64;
        const int e1 = (m >= 2) ? (int)cur[1] : 0;
        words.push_back((uint32_t)ks | ((uint32_t)ncur << 5) | ((uint32_t)nnext << 9) |
                        ((uint32_t)e1 << 13) | ((uint32_t)ends << 21) | ((uint32_t)nf << 22) |
                        ((uint32_t)last << 26));
        words.push_back((uint32_t)adv_exp);
        tp.n_tiles++;
        tp.mma_fma_per_point += 32 * ks;
        tp.fma_per_point += 32 * ks + 8 + (ends ? nf : 0);
        i += ncur + nnext;
    }
    flush();
    tp.slot_src.resize((size_t)tp.n_chunks * MPX_T_COEF_CAP);
}

template <class T>
int upload(T **dst, const std::vector<T> &src)
{
    if (src.empty()) { *dst = nullptr; return MPX_OK; }
    MPX_CUDA(cudaMalloc((void **)dst, src.size() * sizeof(T)));
    MPX_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return MPX_OK;
}


// slice id and slice-local position of every row for the dimension group [lo, hi]
void slice_rows_of_group(const int64_t *e, int64_t N, int m, int lo, int hi,
                         std::vector<int32_t> &slice_of, std::vector<int32_t> &local_of,
                         std::vector<int32_t> &slice_off, std::vector<int32_t> &slice_rows)
{
    std::vector<int32_t> idx(N);
    std::iota(idx.begin(), idx.end(), 0);
    auto outside_less = [&](int32_t a, int32_t b) {
        for (int i = m - 1; i >= 0; --i) {
            if (i >= lo && i <= hi) continue;
            int64_t x = e[(int64_t)a * m + i], y = e[(int64_t)b * m + i];
            if (x != y) return x < y;
        }
        return false;
    };
    // stable: rows of one slice keep their lexicographic order
    std::stable_sort(idx.begin(), idx.end(), outside_less);
    slice_of.assign(N, 0);
    local_of.assign(N, 0);
    slice_off.clear();
    slice_rows.assign(idx.begin(), idx.end());
    int32_t s = -1;
    for (int64_t t = 0; t < N; ++t) {
        if (t == 0 || outside_less(idx[t - 1], idx[t])) {
            ++s;
            slice_off.push_back((int32_t)t);
        }
        slice_of[idx[t]] = s;
        local_of[idx[t]] = (int32_t)(t - slice_off[s]);
    }
    slice_off.push_back((int32_t)N);
    // Small slices are merged with their successors (they stay independent, they just share a
    // CTA): a CTA should own a few hundred rows so that all its warps find fibers to work on.
#ifdef MPX_TUNING
    static const int merge_rows = getenv("MPX_DDS_MERGE_ROWS") ? atoi(getenv("MPX_DDS_MERGE_ROWS")) : MPX_DDS_MERGE_ROWS;
#else
    constexpr int merge_rows = MPX_DDS_MERGE_ROWS;
#endif
    std::vector<int32_t> merged_off;
    int32_t begin = 0;
    for (size_t t = 0; t + 1 < slice_off.size(); ++t) {
        const int32_t end = slice_off[t + 1];
        const bool last = (t + 2 == slice_off.size());
        const int32_t next_end = last ? end : slice_off[t + 2];
        if (last || end - begin >= merge_rows || next_end - begin > MPX_DDS_MAX_ROWS) {
            merged_off.push_back(begin);
            begin = end;
        }
    }
    merged_off.push_back((int32_t)N);
    slice_off.swap(merged_off);
    for (size_t sb = 0; sb + 1 < slice_off.size(); ++sb)
        for (int32_t t = slice_off[sb]; t < slice_off[sb + 1]; ++t) {
            slice_of[idx[t]] = (int32_t)sb;
            local_of[idx[t]] = t - slice_off[sb];
        }
}

int build_dds_groups(mpx_plan *pl, const int64_t *e, const std::vector<std::vector<int32_t>> &fibers_off,
                     const std::vector<std::vector<int32_t>> &fibers_rows, const double *gp_h)
{
    const int m = pl->m;
    const int64_t N = pl->N;
    std::vector<int32_t> slice_of, local_of, slice_off, slice_rows;
    int hi = m - 1;
    while (hi >= 0) {
        // grow the group downwards while its largest slice still fits shared memory
        int lo = hi;
        slice_rows_of_group(e, N, m, lo, hi, slice_of, local_of, slice_off, slice_rows);
        while (lo > 0) {
            std::vector<int32_t> so, lo_, sf, sr;
            slice_rows_of_group(e, N, m, lo - 1, hi, so, lo_, sf, sr);
            int biggest = 0;
            for (size_t t = 0; t + 1 < sf.size(); ++t) biggest = std::max(biggest, sf[t + 1] - sf[t]);
            if (biggest > MPX_DDS_MAX_ROWS) break;
            --lo;
            slice_of.swap(so); local_of.swap(lo_); slice_off.swap(sf); slice_rows.swap(sr);
        }
        DdsGroup grp;
        grp.lo = lo;
        grp.hi = hi;
        grp.n_slices = (int)slice_off.size() - 1;
        for (int t = 0; t < grp.n_slices; ++t) grp.max_rows = std::max(grp.max_rows, slice_off[t + 1] - slice_off[t]);
        if (grp.max_rows > 65535) return mpx_fail(MPX_ERR_UNSUPPORTED, "fiber too long for the dds kernel");
        int rc;
        if ((rc = upload(&grp.d_slice_off, slice_off)) != MPX_OK) return rc;
        if ((rc = upload(&grp.d_slice_rows, slice_rows)) != MPX_OK) return rc;
        for (int d = hi; d >= lo; --d) {
            const std::vector<int32_t> &off = fibers_off[d], &rows = fibers_rows[d];
            const int nf = (int)off.size() - 1;
            // fibers of dimension d, ordered by the slice they live in
            std::vector<int32_t> forder(nf);
            std::iota(forder.begin(), forder.end(), 0);
            std::stable_sort(forder.begin(), forder.end(), [&](int32_t a, int32_t b) {
                return slice_of[rows[off[a]]] < slice_of[rows[off[b]]];
            });
            std::vector<int32_t> fib0(grp.n_slices + 1, 0), foff;
            std::vector<uint16_t> frows;
            frows.reserve(N);
            int L = 1;
            for (int t = 0; t < nf; ++t) {
                const int f = forder[t];
                fib0[slice_of[rows[off[f]]] + 1]++;
                foff.push_back((int32_t)frows.size());
                for (int32_t q = off[f]; q < off[f + 1]; ++q) frows.push_back((uint16_t)local_of[rows[q]]);
                L = std::max(L, off[f + 1] - off[f]);
            }
            foff.push_back((int32_t)frows.size());
            int max_fibers = 0;
            for (int t = 0; t < grp.n_slices; ++t) {
                max_fibers = std::max(max_fibers, fib0[t + 1]);
                fib0[t + 1] += fib0[t];
            }
            std::vector<double> diff((size_t)L * L, 1.0), recip((size_t)L * L, 1.0);
            for (int l = 0; l < L; ++l)
                for (int r = 0; r < L; ++r)
                    if (r != l) {
                        const double dv = gp_h[(size_t)r * m + d] - gp_h[(size_t)l * m + d];
                        diff[(size_t)l * L + r] = dv;
                        recip[(size_t)l * L + r] = reciprocal_or_zero(dv);
                    }
            DdsGroupDim gd;
            gd.dim = d;
            gd.L = L;
            gd.max_fibers = max_fibers;
            if ((rc = upload(&gd.d_slice_fib0, fib0)) != MPX_OK) return rc;
            if ((rc = upload(&gd.d_fib_off, foff)) != MPX_OK) return rc;
            if ((rc = upload(&gd.d_fib_rows, frows)) != MPX_OK) return rc;
            if ((rc = upload(&gd.d_diff, diff)) != MPX_OK) return rc;
            if ((rc = upload(&gd.d_recip, recip)) != MPX_OK) return rc;
            std::vector<double> gcol(L);
            for (int l = 0; l < L; ++l) gcol[l] = gp_h[(size_t)l * m + d];
            if ((rc = upload(&gd.d_g, gcol)) != MPX_OK) return rc;
            grp.dims.push_back(gd);
        }
        pl->dds_groups.push_back(std::move(grp));
        hi = lo - 1;
    }
    return MPX_OK;
}

}  // namespace

extern "C" int mpx_plan_create(int device, const int64_t *exponents_h, int64_t N, int m,
                               const double *gp_h, int64_t gp_rows, mpx_plan **plan_out)
{
    if (!plan_out) return mpx_fail(MPX_ERR_ARG, "plan_out is null");
    *plan_out = nullptr;
    if (!exponents_h || !gp_h) return mpx_fail(MPX_ERR_ARG, "null input");
    if (N < 1) return mpx_fail(MPX_ERR_ARG, "at least 1 monomial must be given");
    if (m < 1) return mpx_fail(MPX_ERR_ARG, "spatial dimension must be >= 1");
    if (m > MPX_MAX_M)
        return mpx_fail(MPX_ERR_UNSUPPORTED, "spatial dimension > 16 is not supported");
    if (N > INT32_MAX / 2) return mpx_fail(MPX_ERR_UNSUPPORTED, "too many monomials");

    mpx_plan *pl = new mpx_plan();
    std::unique_ptr<mpx_plan, void (*)(mpx_plan *)> guard(pl, mpx_plan_destroy);
    pl->device = device;
    pl->N = N;
    pl->m = m;
    const int64_t *e = exponents_h;
    for (int64_t j = 0; j < N; ++j)
        for (int i = 0; i < m; ++i) {
            int64_t v = e[j * m + i];
            if (v < 0) return mpx_fail(MPX_ERR_ARG, "negative exponent");
            if (v > 255) return mpx_fail(MPX_ERR_UNSUPPORTED, "exponent > 255 is not supported");
            if (v > pl->max_exp[i]) pl->max_exp[i] = (int)v;
        }
    int maxmax = 0;
    for (int i = 0; i < m; ++i) {
        // evaluation reads rows < max_exp[i] (jit_compiled_utils.py:102-112)
        if (pl->max_exp[i] > gp_rows)
            return mpx_fail(MPX_ERR_ARG, "generating_points has fewer rows than the largest exponent");
        maxmax = std::max(maxmax, pl->max_exp[i]);
    }
    // beyond 64 the dimension-0 table no longer fits registers: such plans evaluate with the
    // shared-memory-table kernel (mpx_eval_mma.cu) for every k
    pl->tree_ok = pl->max_exp[0] + 1 <= MPX_MAX_L0 - 1;
    pl->gp_rows = (int)std::min<int64_t>(gp_rows, maxmax + 1);

    // lexicographic order (identity for a MultiIndexSet, which is always sorted)
    std::vector<int64_t> order(N);
    std::iota(order.begin(), order.end(), 0);
    LexLess less{e, m};
    bool sorted_unique = true;
    for (int64_t j = 1; j < N && sorted_unique; ++j) sorted_unique = less(j - 1, j);
    if (!sorted_unique) std::stable_sort(order.begin(), order.end(), less);

    // completeness (downward closed): every row with e_d > 0 has its predecessor along d
    bool complete = sorted_unique;
    for (int i = 0; i < m && complete; ++i)
        if (pl->max_exp[i] >= gp_rows) complete = false;    // dds reads row max_exp too
    std::vector<int64_t> key(m);
    for (int64_t j = 0; j < N && complete; ++j)
        for (int i = 0; i < m && complete; ++i)
            if (e[j * m + i] > 0) {
                std::copy(e + j * m, e + j * m + m, key.begin());
                key[i] -= 1;
                if (find_sorted(e, order, m, key.data()) < 0) complete = false;
            }
    pl->complete = complete;

    std::string err;
    int rc = MPX_OK;
    if (pl->tree_ok) {
        rc = compile_program(e, N, m, order, mpx_l0p_for(pl->max_exp[0]), pl->prog, err);
        if (rc != MPX_OK) return mpx_fail(rc, err);
    }
    if (complete && m >= 2 && m <= 8 && pl->max_exp[0] + 1 <= 68 && mpx_test_tmma != 2) {
        compile_tile_program(e, N, m, pl->tprog);
        // Worth it only for long dimension-0 tables (more than 5 k-steps, n > 19) whose tiles
        // carry almost no zero padding -- in practice full tensor grids (lp = inf).  Measured
        // on the B200: cfg5 (9 % more FMAs than the DFMA tree kernel) 1.18e13 vs 9.9e12 pm/s;
        // m=3, n=24, lp=2 (about 25 % more) 5.5e12 vs 9.6e12; cfg2 (about 40 % more, short
        // tiles) 6.8e12 vs 9.3e12.  The test hook mpx_debug_set_tmma(1) takes it whenever it was
        // built, (2) never builds it.
        // (the threshold was tuned on FMA counts that included one more per leaf)
        const int64_t tree_fma = pl->tree_ok ? pl->prog.fma_per_point + pl->prog.n_leaves : 0;
        pl->tprog.usable = mpx_test_tmma == 1 || !pl->tree_ok ||
                           (pl->tprog.ks_max > 5 && 100 * pl->tprog.fma_per_point <= 110 * tree_fma);
    }
    // upper-level exponents of the first leaf (0 for every downward-closed set)
    for (int i = 2; i < m; ++i) {
        pl->start_exp[i] = (int)e[order[0] * m + i];
        pl->prog.other_fp64_per_point += 2 * pl->start_exp[i];
    }
    // table building cost per point: one DADD + one DMUL per table entry
    for (int i = 0; i < m; ++i) pl->prog.other_fp64_per_point += 2 * pl->max_exp[i];

    DeviceGuard dg(device);
    if (!dg.ok) return mpx_fail(MPX_ERR_CUDA, "cannot select CUDA device (no GPU or bad index)");

    if ((rc = upload(&pl->d_desc, pl->prog.desc)) != MPX_OK) return rc;
    if ((rc = upload(&pl->d_slot_src, pl->prog.slot_src)) != MPX_OK) return rc;
    {
        // the used record words as a kernel parameter (mpx_eval_kernel.cuh), if they fit and
        // every record is regular: the uniform kernels index the stream by record number
        const int nc = pl->prog.n_chunks;
        size_t total = 0;
        for (int c = 0; c < nc; ++c) total += MPX_DESC_HDR + pl->prog.desc[(size_t)c * MPX_DESC_WORDS];
        // (no uniform-control instantiation exists for one point per thread: tables of 65 entries)
        bool regular = pl->tree_ok && mpx_l0p_for(pl->max_exp[0]) < 66;
        const int rec_words = 1 + mpx_count_words(mpx_l0p_for(pl->max_exp[0]));
        for (int c = 0; c < nc && regular; ++c) {
            const uint32_t *src = pl->prog.desc.data() + (size_t)c * MPX_DESC_WORDS;
            if (src[0] % rec_words) regular = false;
            for (uint32_t i = 0; i + rec_words <= src[0] && regular; i += rec_words)
                if (src[MPX_DESC_HDR + i] & MPX_D_IRREG) regular = false;
        }
        if (regular && nc <= MPX_U_CHUNKS && total <= (size_t)MPX_U_WORDS) {
            pl->uprog.resize(1);
            MpxUniformProgram &u = pl->uprog[0];
            memset(&u, 0, sizeof(u));
            uint32_t at = 0;
            for (int c = 0; c < nc; ++c) {
                const uint32_t *src = pl->prog.desc.data() + (size_t)c * MPX_DESC_WORDS;
                const uint32_t n = MPX_DESC_HDR + src[0];
                u.off[c] = at;
                std::copy(src, src + n, u.w + at);
                at += n;
            }
            u.off[nc] = at;
        }
    }
    if (pl->tprog.usable) {
        if ((rc = upload(&pl->d_tdesc, pl->tprog.desc)) != MPX_OK) return rc;
        if ((rc = upload(&pl->d_tslot_src, pl->tprog.slot_src)) != MPX_OK) return rc;
    }
    std::vector<double> gpt((size_t)m * pl->gp_rows);
    for (int i = 0; i < m; ++i)
        for (int j = 0; j < pl->gp_rows; ++j) gpt[(size_t)i * pl->gp_rows + j] = gp_h[(size_t)j * m + i];
    if ((rc = upload(&pl->d_gp, gpt)) != MPX_OK) return rc;
    if (m == 1) {
        // Horner kernel (mpx_eval_1d.cu): where each exponent's coefficient sits; duplicates
        // (the reference would add them up) leave the plan on the tree kernel
        std::vector<int32_t> row((size_t)pl->max_exp[0] + 1, -1);
        bool distinct = N <= 0x7fffffffLL;
        for (int64_t j = 0; j < N && distinct; ++j) {
            int32_t &slot = row[(size_t)e[j]];
            if (slot >= 0) distinct = false;
            slot = (int32_t)j;
        }
        if (distinct && (rc = upload(&pl->d_row1d, row)) != MPX_OK) return rc;
    }
    std::vector<uint8_t> e8((size_t)N * m);
    for (size_t t = 0; t < e8.size(); ++t) e8[t] = (uint8_t)e[t];
    if ((rc = upload(&pl->d_exp8, e8)) != MPX_OK) return rc;
    if (m <= 8) {
        pl->mma_chunks = (int)((N + 31) / 32);
        std::vector<uint64_t> e64((size_t)pl->mma_chunks * 32, 0ull);
        for (int64_t j = 0; j < N; ++j)
            for (int i = 0; i < m; ++i) e64[j] |= (uint64_t)e[j * m + i] << (8 * i);
        if ((rc = upload(&pl->d_exp64, e64)) != MPX_OK) return rc;
    }
    std::vector<int> mxv(pl->max_exp, pl->max_exp + m);
    if ((rc = upload(&pl->d_max_exp, mxv)) != MPX_OK) return rc;

    // fiber tables for the divided differences
    std::vector<std::vector<int32_t>> fibers_off(m), fibers_rows(m);
    if (complete) {
        pl->dds.resize(m);
        std::vector<int64_t> succ(N);
        for (int d = 0; d < m; ++d) {
            for (int64_t j = 0; j < N; ++j) {
                std::copy(e + j * m, e + j * m + m, key.begin());
                key[d] += 1;
                succ[j] = (key[d] <= pl->max_exp[d]) ? find_sorted(e, order, m, key.data()) : -1;
            }
            std::vector<int32_t> off, rows;
            rows.reserve(N);
            int max_len = 0;
            for (int64_t j0 = 0; j0 < N; ++j0) {
                if (e[j0 * m + d] != 0) continue;
                off.push_back((int32_t)rows.size());
                int len = 0;
                for (int64_t j = j0; j >= 0; j = succ[j]) { rows.push_back((int32_t)j); ++len; }
                max_len = std::max(max_len, len);
                pl->dds_updates += (int64_t)len * (len - 1) / 2;
            }
            off.push_back((int32_t)rows.size());
            DdsDim &dd = pl->dds[d];
            dd.n_fibers = (int)off.size() - 1;
            dd.max_len = max_len;
            std::vector<double> diff((size_t)max_len * max_len, 1.0);
            for (int l = 0; l < max_len; ++l)
                for (int r = 0; r < max_len; ++r)
                    if (r != l)   // gp[r,d] - gp[l,d], as src/minterpy/dds.py:673-679 / :633
                        diff[(size_t)l * max_len + r] = gp_h[(size_t)r * m + d] - gp_h[(size_t)l * m + d];
            if ((rc = upload(&dd.d_off, off)) != MPX_OK) return rc;
            if ((rc = upload(&dd.d_rows, rows)) != MPX_OK) return rc;
            if ((rc = upload(&dd.d_diff, diff)) != MPX_OK) return rc;
            {
                // work list: fibers with something to do, longest first (the tail of the launch
                // is then made of short fibers)
                const int nfib = dd.n_fibers;
                std::vector<int> byl(nfib);
                std::iota(byl.begin(), byl.end(), 0);
                std::stable_sort(byl.begin(), byl.end(), [&](int a, int b) {
                    return off[a + 1] - off[a] > off[b + 1] - off[b];
                });
                std::vector<int32_t> woff, wrows;
                for (int f : byl) {
                    const int len = off[f + 1] - off[f];
                    if (len < 2) break;
                    woff.push_back((int32_t)wrows.size());
                    wrows.insert(wrows.end(), rows.begin() + off[f], rows.begin() + off[f + 1]);
                }
                dd.n_work = (int)woff.size();
                woff.push_back((int32_t)wrows.size());
                if (wrows.empty()) wrows.push_back(0);
                std::vector<double> dy((size_t)2 * max_len * max_len, 1.0), gcol(max_len);
                for (int l = 0; l < max_len; ++l) {
                    gcol[l] = gp_h[(size_t)l * m + d];
                    for (int r = 0; r < max_len; ++r)
                        if (r != l) {
                            const double dv = diff[(size_t)l * max_len + r];
                            dy[2 * ((size_t)l * max_len + r)] = dv;
                            dy[2 * ((size_t)l * max_len + r) + 1] = reciprocal_or_zero(dv);
                        }
                }
                if (max_len <= 40) {
                    std::vector<double> dy2v(dy);
                    std::vector<unsigned long long> fmask(max_len, 0ull);
                    dd.cols_ok = true;
                    for (int l = 0; l < max_len; ++l)
                        for (int r = l + 1; r < max_len; ++r) {
                            const size_t at = 2 * ((size_t)l * max_len + r);
                            if (dy[at + 1] != 0.0) continue;             // ordinary divisor
                            double pw = 0.0;
                            if (!all_ones_divisor(dy[at], &pw)) { dd.cols_ok = false; continue; }
                            dy2v[at] = pw;                               // +-2^j
                            dy2v[at + 1] = 1.0 / pw;                     // +-2^-j, exact
                            fmask[l] |= 1ull << r;
                        }
                    if (dd.cols_ok) {
                        if ((rc = upload(&dd.d_dy2, dy2v)) != MPX_OK) return rc;
                        if ((rc = upload(&dd.d_fmask, fmask)) != MPX_OK) return rc;
                        // rows at a fixed stride (slots behind a fiber's end repeat its last row)
                        std::vector<int32_t> wpad((size_t)dd.n_work * max_len + 64, 0);
                        for (int f = 0; f < dd.n_work; ++f) {
                            const int len = woff[f + 1] - woff[f];
                            for (int r = 0; r < max_len; ++r)
                                wpad[(size_t)f * max_len + r] = wrows[woff[f] + (r < len ? r : len - 1)];
                        }
                        if ((rc = upload(&dd.d_wpad, wpad)) != MPX_OK) return rc;
                    }
                }
                if (max_len <= MPX_DDS_REG_LEN) {
                    dd.h_pairs.assign((size_t)MPX_DDS_REG_LEN * (MPX_DDS_REG_LEN - 1), 1.0);
                    dd.all_fast = true;
                    for (int r = 1; r < max_len; ++r)
                        for (int l = 0; l < r; ++l) {
                            const size_t at = 2 * ((size_t)r * (r - 1) / 2 + l);
                            dd.h_pairs[at] = dy[2 * ((size_t)l * max_len + r)];
                            dd.h_pairs[at + 1] = dy[2 * ((size_t)l * max_len + r) + 1];
                            dd.all_fast = dd.all_fast && dd.h_pairs[at + 1] != 0.0;
                        }
                }
                if ((rc = upload(&dd.d_woff, woff)) != MPX_OK) return rc;
                if ((rc = upload(&dd.d_wrows, wrows)) != MPX_OK) return rc;
                if ((rc = upload(&dd.d_dy, dy)) != MPX_OK) return rc;
                if ((rc = upload(&dd.d_gcol, gcol)) != MPX_OK) return rc;
            }
            fibers_off[d] = std::move(off);
            fibers_rows[d] = std::move(rows);
        }
    }
    if (complete && (rc = build_dds_groups(pl, e, fibers_off, fibers_rows, gp_h)) != MPX_OK) return rc;
    guard.release();
    *plan_out = pl;
    return MPX_OK;
}

extern "C" void mpx_plan_destroy(mpx_plan *pl)
{
    if (!pl) return;
    DeviceGuard dg(pl->device);
    cudaFree(pl->d_desc);
    cudaFree(pl->d_slot_src);
    cudaFree(pl->d_row1d);
    cudaFree(pl->d_tdesc);
    cudaFree(pl->d_tslot_src);
    cudaFree(pl->d_gp);
    cudaFree(pl->d_exp8);
    cudaFree(pl->d_max_exp);
    cudaFree(pl->d_exp64);
    mpx_image_destroy(&pl->image);
    mpx_image_destroy(&pl->host_image);
    mpx_image_destroy(&pl->deriv_tables);
    for (int b = 0; b < 2; ++b) {
        cudaFree(pl->h_x[b]);
        cudaFree(pl->h_o[b]);
        if (pl->p_x[b]) cudaFreeHost(pl->p_x[b]);
        if (pl->p_o[b]) cudaFreeHost(pl->p_o[b]);
        if (pl->h_stream[b]) cudaStreamDestroy(pl->h_stream[b]);
        if (pl->h_done[b]) cudaEventDestroy(pl->h_done[b]);
        if (pl->h_up[b]) cudaEventDestroy(pl->h_up[b]);
    }
    cudaFree(pl->h_c);
    cudaFree(pl->h_v);
    if (pl->h_event) cudaEventDestroy(pl->h_event);
    for (DdsGroup &g : pl->dds_groups) {
        cudaFree(g.d_slice_off);
        cudaFree(g.d_slice_rows);
        for (DdsGroupDim &gd : g.dims) {
            cudaFree(gd.d_slice_fib0);
            cudaFree(gd.d_fib_off);
            cudaFree(gd.d_fib_rows);
            cudaFree(gd.d_diff);
            cudaFree(gd.d_recip);
            cudaFree(gd.d_g);
        }
    }
    for (DdsDim &dd : pl->dds) {
        cudaFree(dd.d_off);
        cudaFree(dd.d_rows);
        cudaFree(dd.d_diff);
        cudaFree(dd.d_woff);
        cudaFree(dd.d_wrows);
        cudaFree(dd.d_wpad);
        cudaFree(dd.d_dy);
        cudaFree(dd.d_dy2);
        cudaFree(dd.d_fmask);
        cudaFree(dd.d_gcol);
    }
    delete pl;
}

extern "C" int mpx_plan_info(const mpx_plan *pl, int what, int64_t *value_out)
{
    if (!pl || !value_out) return mpx_fail(MPX_ERR_ARG, "null argument");
    switch (what) {
    case MPX_INFO_N: *value_out = pl->N; break;
    case MPX_INFO_M: *value_out = pl->m; break;
    case MPX_INFO_COMPLETE: *value_out = pl->complete ? 1 : 0; break;
    // (Horner kernel, m = 1: one FMA and one subtraction per exponent up to the largest)
    case MPX_INFO_EVAL_FMA: *value_out = mpx_1d_eligible(pl, 1) ? pl->max_exp[0] : pl->prog.fma_per_point; break;
    case MPX_INFO_EVAL_OTHER_FP64:
        *value_out = mpx_1d_eligible(pl, 1) ? pl->max_exp[0] : pl->prog.other_fp64_per_point;
        break;
    case MPX_INFO_DDS_UPDATES: *value_out = pl->dds_updates; break;
    case MPX_INFO_LEAVES: *value_out = pl->prog.n_leaves; break;
    case MPX_INFO_DEVICE: *value_out = pl->device; break;
    case MPX_INFO_SCRATCH_BYTES: *value_out = (int64_t)pl->image.bytes; break;
    case MPX_INFO_CHUNKS: *value_out = pl->prog.n_chunks; break;
    case MPX_INFO_SMALLK_KERNEL:
        *value_out = mpx_1d_eligible(pl, 1) ? 3 : mpx_tmma_eligible(pl, 1) ? 1 : (pl->tree_ok ? 0 : 2);
        break;
    case MPX_INFO_TMMA_FMA: *value_out = pl->tprog.fma_per_point; break;
    case MPX_INFO_UNIFORM_PROGRAM: *value_out = pl->uprog.empty() ? 0 : 1; break;
    case MPX_INFO_MMA_MIN_K:
        *value_out = mpx_1d_eligible(pl, 1) ? 0 : (pl->m <= 8 && pl->d_exp64) ? (pl->tree_ok ? mpx_mma_min_k(pl) : 1) : 0;
        break;
    default: return mpx_fail(MPX_ERR_ARG, "unknown info key");
    }
    return MPX_OK;
}

bool mpx_smem_raised(const void *fn, int device, bool mark)
{
    static std::mutex mu;
    static std::set<std::pair<const void *, int>> done;
    std::lock_guard<std::mutex> lk(mu);
    if (mark) {
        done.insert({fn, device});
        return true;
    }
    return done.count({fn, device}) != 0;
}

int mpx_image_acquire(MpxImage *img, size_t bytes, cudaStream_t stream)
{
    if (bytes > img->bytes) {
        if (img->fixed)
            return mpx_fail(MPX_ERR_UNSUPPORTED, "packed coefficients exceed the fixed scratch array");
        // cudaFree waits for all work on the device, readers of the old image included
        if (img->ptr) MPX_CUDA(cudaFree(img->ptr));
        img->ptr = nullptr;
        img->bytes = 0;
        img->in_use = false;
        const size_t want = std::max(bytes, (size_t)1 << 20);
        MPX_CUDA(cudaMalloc((void **)&img->ptr, want));
        img->bytes = want;
    }
    if (img->in_use && img->last_stream != stream) {
        // a capturing stream must not wait on an event from outside its graph; such callers
        // order their streams themselves (as the header asks)
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        MPX_CUDA(cudaStreamIsCapturing(stream, &cap));
        if (cap == cudaStreamCaptureStatusNone) MPX_CUDA(cudaStreamWaitEvent(stream, img->busy, 0));
    }
    return MPX_OK;
}

int mpx_image_release(MpxImage *img, cudaStream_t stream)
{
    if (img->fixed) return MPX_OK;      // fixed images: the caller serialises (documented)
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    MPX_CUDA(cudaStreamIsCapturing(stream, &cap));
    if (cap != cudaStreamCaptureStatusNone) return MPX_OK;
    if (!img->busy) MPX_CUDA(cudaEventCreateWithFlags(&img->busy, cudaEventDisableTiming));
    MPX_CUDA(cudaEventRecord(img->busy, stream));
    img->last_stream = stream;
    img->in_use = true;
    return MPX_OK;
}

void mpx_image_destroy(MpxImage *img)
{
    if (!img->fixed) cudaFree(img->ptr);
    if (img->busy) cudaEventDestroy(img->busy);
    *img = MpxImage();
}

// Host-only view of the compiled evaluation program, for tests of the index-set compiler
// (no device needed).  desc_out / slot_out may be null to query the sizes.
extern "C" int mpx_debug_program(const int64_t *exponents_h, int64_t N, int m, uint32_t *desc_out,
                                 int32_t *slot_out, int64_t *n_chunks_out, int64_t *fma_out)
{
    if (!exponents_h || N < 1 || m < 1 || m > MPX_MAX_M) return mpx_fail(MPX_ERR_ARG, "bad argument");
    for (int64_t t = 0; t < N * m; ++t)
        if (exponents_h[t] < 0 || exponents_h[t] > 255) return mpx_fail(MPX_ERR_UNSUPPORTED, "exponent out of range");
    std::vector<int64_t> order(N);
    std::iota(order.begin(), order.end(), 0);
    LexLess less{exponents_h, m};
    std::stable_sort(order.begin(), order.end(), less);
    EvalProgram pg;
    std::string err;
    int max0 = 0;
    for (int64_t j = 0; j < N; ++j) max0 = std::max<int>(max0, (int)exponents_h[j * m]);
    if (max0 > MPX_MAX_L0 - 2) return mpx_fail(MPX_ERR_UNSUPPORTED, "dimension-0 exponent too large");
    int rc = compile_program(exponents_h, N, m, order, mpx_l0p_for(max0), pg, err);
    if (rc != MPX_OK) return mpx_fail(rc, err);
    if (n_chunks_out) *n_chunks_out = pg.n_chunks;
    if (fma_out) *fma_out = pg.fma_per_point;
    if (desc_out) std::copy(pg.desc.begin(), pg.desc.end(), desc_out);
    if (slot_out) std::copy(pg.slot_src.begin(), pg.slot_src.end(), slot_out);
    return MPX_OK;
}

// Host-only view of the tile program (tensor-core tree kernel), for the CPU tests.
extern "C" int mpx_debug_tile_program(const int64_t *exponents_h, int64_t N, int m, uint32_t *desc_out,
                                      int32_t *slot_out, int64_t *n_chunks_out, int64_t *fma_out)
{
    if (!exponents_h || N < 1 || m < 2 || m > MPX_MAX_M) return mpx_fail(MPX_ERR_ARG, "bad argument");
    TileProgram tp;
    compile_tile_program(exponents_h, N, m, tp);
    if (n_chunks_out) *n_chunks_out = tp.n_chunks;
    if (fma_out) *fma_out = tp.fma_per_point;
    if (desc_out) std::copy(tp.desc.begin(), tp.desc.end(), desc_out);
    if (slot_out) std::copy(tp.slot_src.begin(), tp.slot_src.end(), slot_out);
    return MPX_OK;
}
