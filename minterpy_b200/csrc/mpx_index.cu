// Host-side index-set queries of the C ABI (no device work, usable without a GPU).
//   replaces: is_lexicographically_complete  src/minterpy/core/utils.py:234-262
//             (MultiIndexSet.is_complete, core/multi_index.py:120-131), which walks every
//             row's "one smaller" neighbours with a linear search, O(N^2 m) in Python --
//             25 s at cfg2 and minutes at cfg4 before any kernel runs.
// The same predicate is one merge per dimension here: for rows in lexicographic order the
// neighbours e - u_d (rows with e_d > 0) come out in lexicographic order too, so a single
// forward scan of the set finds them all -- O(N m^2) comparisons of 8-byte words.
#include <vector>

#include "mpx_common.h"

namespace {

// -1 / 0 / +1 for row a against row b with component d of b lowered by `dec`
inline int lex_cmp(const int64_t *a, const int64_t *b, int m, int d, int64_t dec)
{
    for (int i = m - 1; i >= 0; --i) {
        const int64_t x = a[i], y = b[i] - (i == d ? dec : 0);
        if (x != y) return x < y ? -1 : 1;
    }
    return 0;
}

}  // namespace

extern "C" int mpx_index_is_sorted(const int64_t *exponents_h, int64_t N, int m, int *sorted_out)
{
    if (!exponents_h || !sorted_out) return mpx_fail(MPX_ERR_ARG, "null argument");
    if (N < 0 || m < 1) return mpx_fail(MPX_ERR_ARG, "bad shape");
    int ok = 1;
    for (int64_t j = 1; j < N && ok; ++j)
        ok = lex_cmp(exponents_h + (j - 1) * m, exponents_h + j * m, m, -1, 0) < 0;
    *sorted_out = ok;
    return MPX_OK;
}

extern "C" int mpx_index_is_complete(const int64_t *exponents_h, int64_t N, int m,
                                     int *complete_out)
{
    if (!exponents_h || !complete_out) return mpx_fail(MPX_ERR_ARG, "null argument");
    if (N < 1) return mpx_fail(MPX_ERR_ARG, "at least 1 monomial must be given");
    if (m < 1) return mpx_fail(MPX_ERR_ARG, "spatial dimension must be >= 1");
    const int64_t *e = exponents_h;
    for (int64_t j = 0; j < N * m; ++j)
        if (e[j] < 0) return mpx_fail(MPX_ERR_ARG, "negative exponent");
    int sorted = 0;
    mpx_index_is_sorted(e, N, m, &sorted);
    if (!sorted)
        return mpx_fail(MPX_ERR_ARG, "the exponents must be lexicographically sorted (last "
                                     "column most significant) without duplicates");
    *complete_out = 0;
    for (int d = 0; d < m; ++d) {
        int64_t s = 0;                       // scan position in the set
        for (int64_t j = 0; j < N; ++j) {
            const int64_t *row = e + j * m;
            if (row[d] == 0) continue;
            // the neighbour precedes row j; neighbours arrive in increasing order, so the
            // scan never has to move backwards
            int c = -1;
            while (s < j && (c = lex_cmp(e + s * m, row, m, d, 1)) < 0) ++s;
            if (s >= j || c != 0) return MPX_OK;      // a hole: not complete
        }
    }
    *complete_out = 1;
    return MPX_OK;
}
