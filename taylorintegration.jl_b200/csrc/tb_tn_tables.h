/*
 * tb_tn_tables.h -- index tables of truncated multivariate polynomials (TaylorN),
 * shared by the library host code and the CPU oracle (tables only, no arithmetic).
 *
 * A TaylorN in n variables of maximal degree d is stored flat: homogeneous parts by
 * ascending degree, and inside one degree the monomials x1^e1 ... xn^en in
 * lexicographically DESCENDING exponent order (x1^2, x1 x2, x2^2 for n = 2), the
 * flat layout the reference serialises (src/integrator/taylorsolution.jl:596-609;
 * the within-degree order is TaylorSeries' own and is recalled, not read: SURVEY
 * App. B / hard part 7).  Number of coefficients: C(n+d, d).
 *
 * Products are evaluated per OUTPUT monomial m as a sum over the pairs (ia, ib) with
 * mono[ia] + mono[ib] = mono[m], in ascending ia.  That is the order in which
 * TaylorSeries' scatter loops (for i in degrees, for na, for nb: c[pos] += a[na]*b[nb])
 * touch slot m, so a gather over this list reproduces their rounding.
 */
#ifndef TB_TN_TABLES_H
#define TB_TN_TABLES_H

#include <cstdint>
#include <map>
#include <vector>

#if defined(__GNUC__) && !defined(__clang__)
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wstringop-overflow"
#pragma GCC diagnostic ignored "-Warray-bounds"
#endif

namespace tbtn {

struct Tables {
    int n = 0, d = 0, ncoef = 0;
    std::vector<int> deg_off;            /* [d+2] first flat index of each degree             */
    std::vector<uint8_t> expo;           /* [ncoef*n] exponents                               */
    std::vector<uint8_t> deg;            /* [ncoef] total degree                              */
    std::vector<int32_t> pair_off;       /* [ncoef+1] CSR offsets into pair_a / pair_b        */
    std::vector<uint16_t> pair_a, pair_b;
};

inline void enum_monos(int n, int g, std::vector<uint8_t> &cur, int pos, std::vector<std::vector<uint8_t>> &out) {
    if (pos == n - 1) { cur[pos] = (uint8_t)g; out.push_back(cur); return; }
    for (int e = g; e >= 0; e--) { cur[pos] = (uint8_t)e; enum_monos(n, g - e, cur, pos + 1, out); }
}

inline Tables make_tables(int n, int d) {
    Tables T;
    T.n = n; T.d = d;
    std::vector<std::vector<uint8_t>> monos;
    T.deg_off.push_back(0);
    for (int g = 0; g <= d; g++) {
        std::vector<uint8_t> cur(n, 0);
        enum_monos(n, g, cur, 0, monos);
        T.deg_off.push_back((int)monos.size());
    }
    T.ncoef = (int)monos.size();
    std::map<std::vector<uint8_t>, int> index;
    for (int i = 0; i < T.ncoef; i++) {
        index[monos[i]] = i;
        int s = 0;
        for (int v = 0; v < n; v++) { T.expo.push_back(monos[i][v]); s += monos[i][v]; }
        T.deg.push_back((uint8_t)s);
    }
    std::vector<std::vector<std::pair<uint16_t, uint16_t>>> lists(T.ncoef);
    std::vector<uint8_t> sum(n);
    for (int ia = 0; ia < T.ncoef; ia++)
        for (int ib = 0; ib < T.ncoef; ib++) {
            if (T.deg[ia] + T.deg[ib] > d) break; /* ib ascending => degree ascending */
            for (int v = 0; v < n; v++) sum[v] = (uint8_t)(monos[ia][v] + monos[ib][v]);
            lists[index[sum]].push_back({(uint16_t)ia, (uint16_t)ib});
        }
    T.pair_off.push_back(0);
    for (int m = 0; m < T.ncoef; m++) {
        for (auto &p : lists[m]) { T.pair_a.push_back(p.first); T.pair_b.push_back(p.second); }
        T.pair_off.push_back((int32_t)T.pair_a.size());
    }
    return T;
}

} // namespace tbtn

#if defined(__GNUC__) && !defined(__clang__)
#pragma GCC diagnostic pop
#endif

#endif
