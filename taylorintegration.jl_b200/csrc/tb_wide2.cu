/*
 * tb_wide2.cu -- large-state ODEs with the jets in SHARED MEMORY (the many-spin system of
 * benchmark/manyspin_benchmarks.jl:10-41 at N = 100 spins: BASELINE.json config 4, "large-state jetcoeffs! in shared
 * memory").  Same drivers as tb_wide.cu (taylorinteg! src/integrator/taylorinteg.jl:329-381, 582-642; taylorstep!
 * src/integrator/taylorstep.jl:59-82; stepsize src/integrator/stepsize.jl:58-132; solcoeff! src/TaylorIntegration.jl:
 * 33-40) and the same recurrences in the oracle's order, so parity mode is bit-identical to tb_wide.cu and the oracle.
 *
 * What is different from tb_wide.cu (one thread per tape entry per dependency level, jets in a global workspace,
 * ncu: 21 stall cycles per issue on L1/L2 loads, 13 on barriers, FP64 pipe 1.7 % busy):
 *   - CLUSTERS.  Entries of one order that depend on each other (through a series computed in that same order) form a
 *     cluster; clusters only read each other's STATE jets, which solcoeff! finished one order earlier.  One thread
 *     runs a whole cluster in tape order, so an order needs ONE barrier (many-spin: cluster i = everything of spin i,
 *     100 clusters: the 100-term sum over the spins, two Cauchy products, a few scalings, three solcoeff!).
 *   - PERSISTENT vs TRANSIENT series.  Only series read at another index than the current order (Cauchy-product
 *     operands, state jets) are stored whole: [k][series] in shared memory, series numbered role-major so that the
 *     threads of a warp touch consecutive words (no bank conflicts); the coupling sum reads x_a[k] as a broadcast.
 *     Every other series is a scalar of the current order in a [slot][thread] scratch.  A negated series that only
 *     feeds Cauchy products is not stored: the product negates its operand (exact).  Many-spin 100: 400 series x 26
 *     doubles = 83 KB per CTA, two CTAs per SM.
 *   - GROUP PROGRAMS.  Isomorphic clusters share ONE copy of their entries (operands shift by the cluster's rank),
 *     held in shared memory with the per-cluster literals: every lane reads the same entry (broadcast).
 *   - coupling constants come from global memory term-major ([term][cluster]: coalesced, L1/L2 resident).
 * Bound: FP64 pipe; the coupling sum is one dependent DADD chain per thread (parity order).
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <numeric>
#include <string>
#include <vector>

#include "tb_internal.h"
#include "tb_math.h"
#include "tb_wide_prog.h"

namespace tb {
void count_launch(long long n);
void set_last_kernel_ms(double ms);
}

namespace tbw {

#define TB_CUDA2(call)                                                         \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            tb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
            return TB_ERR_CUDA;                                                \
        }                                                                      \
    } while (0)

#include "tb_wide2_dev.cuh"


/* ---- host: fused program -> clusters, persistent / transient placement ------------------------------------- */
struct Wide2 {
    int order = 0, mode = 0, dof = 0, npar = 0, NS = 0, ncl = 0, maxlen = 0, ntr = 0, tsid = -1, threads = 0, ctas = 1;
    size_t smem = 0;
    bool lc_smem = false, g_smem = false;
    std::vector<C2Op> ops;
    std::vector<double> lc_coef;
    std::vector<int> lc_ser, xsid, cl_len;
    std::vector<C2Op> gops;      /* group programs (empty: per-cluster entries) */
    std::vector<double> gnum;    /* literal of entry j for cluster rank r */
    std::vector<int> g_first, g_count, g_off, g_len;
    C2Op *d_gops = nullptr;
    double *d_gnum = nullptr;
    /* NVRTC-specialised build of the group programs (wide2_generate); empty cubin: the interpreter runs */
    std::vector<char> gen_cubin;
    std::string gen_source, gen_log;
    size_t gen_smem = 0;
    int gen_ctas = 1, gen_regs = 0, gen_local = 0;
    bool gen_lc_smem = false;
    std::map<int, std::pair<void *, void *>> gen_loaded; /* device -> (module, function) */
    std::mutex gen_mu;                                   /* guards gen_loaded (calls from several host threads) */
    int device = -1;
    C2Op *d_ops = nullptr;
    double *d_lcc = nullptr;
    int *d_lcs = nullptr, *d_xs = nullptr, *d_len = nullptr;
};

static bool cauchy_op(const WOp &o, int which /*0 a, 1 b, 2 res, 3 aux*/) {
    const bool sa = o.a < 0, sb = o.b < 0;
    int op = o.op;
    if (op == TB_OP_POW) {
        if (o.imm == 0.0 || o.imm == 1.0) return false;
        if (o.imm == 2.0) op = TB_OP_SQR;
        else if (o.imm == 0.5) op = TB_OP_SQRT;
    }
    switch (op) {
        case TB_OP_MUL: return !sa && !sb && which <= 1;
        case TB_OP_SQR: return which == 0;
        case TB_OP_DIV: return !sb && (which == 1 || which == 2);
        case TB_OP_SQRT: return which == 2;
        case TB_OP_POW: case TB_OP_EXP: case TB_OP_LOG: return which == 0 || which == 2;
        case TB_OP_SINCOS: return which == 0 || which == 2 || which == 3;
        default: return false;
    }
}

static std::string wide2_generate(const Wide2 *w, bool lc_smem, int ctas, int lcb, int lcreg);

Wide2 *wide2_build(const Program &p, std::string &why) {
    const int n = (int)p.ops.size(), nser = p.nser;
    if (p.npar > 8) { why = "more than 8 parameters"; return nullptr; }
    std::vector<int> producer(nser, -1), is_state(nser, 0), is_dx(nser, 0);
    for (int c = 0; c < p.dof; c++) { is_state[p.xser[c]] = 1; is_dx[p.dxser[c]] = 1; }
    for (int j = 0; j < n; j++) {
        producer[p.ops[j].res] = j;
        if (p.ops[j].op == TB_OP_SINCOS && p.ops[j].aux >= 0) producer[p.ops[j].aux] = j;
    }
    /* uses of every series: total, and as operand of a series x series product */
    std::vector<int> uses(nser, 0), mul_uses(nser, 0);
    for (int j = 0; j < n; j++) {
        const WOp &o = p.ops[j];
        const bool cm = o.op == TB_OP_MUL && o.a >= 0 && o.b >= 0;
        if (o.a >= 0) { uses[o.a]++; if (cm) mul_uses[o.a]++; }
        if (o.b >= 0) { uses[o.b]++; if (cm) mul_uses[o.b]++; }
        if (o.op == W_LINCOMB) for (int t = 0; t < o.lc_n; t++) uses[p.lc_ser[o.lc_off + t]]++;
    }
    /* a negation that only feeds Cauchy products is folded into them (-(y) * s: the operand is negated on the fly) */
    std::vector<int> alias(nser, -1), alias_neg(nser, 0), dead(n, 0);
    for (int j = 0; j < n; j++) {
        const WOp &o = p.ops[j];
        if (o.op != TB_OP_NEG || o.a < 0 || is_dx[o.res] || uses[o.res] == 0 || uses[o.res] != mul_uses[o.res]) continue;
        int src = o.a, neg = 1;
        if (alias[src] >= 0) { neg ^= alias_neg[src]; src = alias[src]; }
        alias[o.res] = src; alias_neg[o.res] = neg; dead[j] = 1;
    }
    auto resolve = [&](int s, int *neg) { *neg = 0; if (s >= 0 && alias[s] >= 0) { *neg = alias_neg[s]; return alias[s]; } return s; };
    /* persistent series: state jets, the time series, operands / results of recurrences */
    std::vector<int> persistent(nser, 0);
    for (int s = 0; s < nser; s++) if (is_state[s]) persistent[s] = 1;
    if (p.tser >= 0) persistent[p.tser] = 1;
    for (int j = 0; j < n; j++) {
        if (dead[j]) continue;
        const WOp &o = p.ops[j];
        int ng;
        if (o.a >= 0 && cauchy_op(o, 0)) persistent[resolve(o.a, &ng)] = 1;
        if (o.b >= 0 && cauchy_op(o, 1)) persistent[resolve(o.b, &ng)] = 1;
        if (cauchy_op(o, 2)) persistent[o.res] = 1;
        if (o.aux >= 0 && cauchy_op(o, 3)) persistent[o.aux] = 1;
    }
    /* clusters: union-find over the entries along same-order dependencies */
    std::vector<int> uf(n);
    std::iota(uf.begin(), uf.end(), 0);
    auto find = [&](int x) { while (uf[x] != x) { uf[x] = uf[uf[x]]; x = uf[x]; } return x; };
    auto link = [&](int j, int s) {
        int ng;
        s = resolve(s, &ng);
        if (s < 0 || producer[s] < 0) return;
        uf[find(j)] = find(producer[s]);
    };
    for (int j = 0; j < n; j++) {
        if (dead[j]) continue;
        const WOp &o = p.ops[j];
        link(j, o.a); link(j, o.b);
        if (o.op == W_LINCOMB) for (int t = 0; t < o.lc_n; t++) link(j, p.lc_ser[o.lc_off + t]);
    }
    std::map<int, int> cl_of_root;
    std::vector<std::vector<int>> cl_ops; /* tape indices, ascending */
    std::vector<int> cl_of_op(n, -1);
    for (int j = 0; j < n; j++) {
        if (dead[j]) continue;
        const int r = find(j);
        auto it = cl_of_root.find(r);
        if (it == cl_of_root.end()) { it = cl_of_root.emplace(r, (int)cl_ops.size()).first; cl_ops.emplace_back(); }
        cl_ops[it->second].push_back(j);
        cl_of_op[j] = it->second;
    }
    int ncl = (int)cl_ops.size();
    int threads_pick = 32;
    std::vector<int> grp_first, grp_count; /* isomorphic clusters, contiguous in the layout */
    if (ncl == 0) { why = "empty program"; return nullptr; }
    /* solcoeff! of component c runs in the cluster that produces dx[c] */
    std::vector<std::vector<int>> cl_sol(ncl);
    for (int c = 0; c < p.dof; c++) {
        const int pj = producer[p.dxser[c]];
        if (pj < 0 || dead[pj]) { why = "dx without a producer"; return nullptr; }
        cl_sol[cl_of_op[pj]].push_back(c);
    }
    /* Thread assignment.  Clusters with the same opcode sequence (the same role in the system: "spin i") are laid out
     * next to each other so that the lanes of a warp run the same entries; the groups are ordered by cost, the CTA gets
     * one thread per member of the most expensive group, and every group starts at a multiple of the CTA size, so that
     * thread t runs the t-th member of EVERY group (many-spin: the coupling cluster of spin t and its small
     * dq[3t-2] = -B y cluster).  Slots that pad a group hold an empty cluster.                                        */
    {
        std::vector<double> cost(ncl, 0.0);
        std::vector<std::string> sig(ncl);
        for (int c = 0; c < ncl; c++) {
            for (int j : cl_ops[c]) {
                const WOp &o = p.ops[j];
                sig[c] += std::to_string(o.op) + (o.a < 0 ? "n" : "s") + (o.b < 0 ? "n" : "s") + ",";
                cost[c] += o.op == W_LINCOMB ? o.lc_n : ((cauchy_op(o, 0) || cauchy_op(o, 1) || cauchy_op(o, 2)) ? p.order / 2.0 : 1.0);
            }
            sig[c] += "|" + std::to_string(cl_sol[c].size());
        }
        std::map<std::string, std::vector<int>> groups;
        for (int c = 0; c < ncl; c++) groups[sig[c]].push_back(c);
        std::vector<std::pair<double, std::string>> order;
        for (auto &kv : groups) {
            double w = 0.0;
            for (int c : kv.second) w = std::max(w, cost[c]);
            order.push_back({-w, kv.first});
        }
        std::sort(order.begin(), order.end());
        const int lead = (int)groups[order[0].second].size();
        threads_pick = std::max(32, std::min(256, ((lead + 31) / 32) * 32));
        std::vector<std::vector<int>> new_ops, new_sol;
        for (auto &og : order) {
            const std::vector<int> &mem = groups[og.second];
            grp_first.push_back((int)new_ops.size());
            grp_count.push_back((int)mem.size());
            for (int c : mem) { new_ops.push_back(cl_ops[c]); new_sol.push_back(cl_sol[c]); }
            if (&og != &order.back())
                while (new_ops.size() % (size_t)threads_pick) { new_ops.emplace_back(); new_sol.emplace_back(); }
        }
        cl_ops.swap(new_ops);
        cl_sol.swap(new_sol);
    }
    ncl = (int)cl_ops.size(); /* with the padding slots */
    /* role-major numbering of the persistent series: role r of the clusters that have it, in layout order */
    std::vector<int> pid(nser, -1), tslot(nser, -1);
    int nroles = 0, ntr = 0;
    std::vector<std::vector<int>> cl_pers(ncl);
    for (int c = 0; c < ncl; c++) {
        int nt = 0;
        auto own = [&](int s) {
            if (s < 0) return;
            if (persistent[s]) { if (std::find(cl_pers[c].begin(), cl_pers[c].end(), s) == cl_pers[c].end()) cl_pers[c].push_back(s); }
            else if (tslot[s] < 0) tslot[s] = nt++;
        };
        for (int comp : cl_sol[c]) own(p.xser[comp]);
        for (int j : cl_ops[c]) { own(p.ops[j].res); if (p.ops[j].op == TB_OP_SINCOS) own(p.ops[j].aux); }
        nroles = std::max(nroles, (int)cl_pers[c].size());
        ntr = std::max(ntr, nt);
    }
    int NS = 0, tsid = -1;
    for (int r = 0; r < nroles; r++)
        for (int c = 0; c < ncl; c++)
            if ((int)cl_pers[c].size() > r) pid[cl_pers[c][r]] = NS++;
    if (p.tser >= 0) { tsid = NS++; pid[p.tser] = tsid; }
    for (int s = 0; s < nser; s++)
        if (persistent[s] && pid[s] < 0) { why = "a stored series has no owner cluster"; return nullptr; }
    Wide2 *w = new Wide2;
    w->order = p.order; w->mode = p.mode; w->dof = p.dof; w->npar = p.npar; w->NS = NS; w->ncl = ncl; w->ntr = std::max(ntr, 1); w->tsid = tsid;
    for (int c = 0; c < p.dof; c++) w->xsid.push_back(pid[p.xser[c]]);
    int maxlen = 0;
    for (int c = 0; c < ncl; c++) {
        w->cl_len.push_back((int)(cl_ops[c].size() + cl_sol[c].size()));
        maxlen = std::max(maxlen, w->cl_len.back());
    }
    w->maxlen = maxlen;
    w->ops.assign((size_t)maxlen * ncl, C2Op{});
    auto loc = [&](int s, int *neg) -> int {
        *neg = 0;
        if (s < 0) return -1;
        s = resolve(s, neg);
        if (persistent[s]) return pid[s];
        return -2 - tslot[s];
    };
    std::string bad;
    /* linear-combination terms: coefficients term-major [t][cluster] per entry position, series ids shared when uniform */
    for (int pos = 0; pos < maxlen; pos++) {
        int longest = 0;
        bool uniform = true;
        const WOp *first = nullptr;
        for (int c = 0; c < ncl; c++) {
            if (pos >= (int)cl_ops[c].size()) continue;
            const WOp &o = p.ops[cl_ops[c][pos]];
            if (o.op != W_LINCOMB) continue;
            longest = std::max(longest, o.lc_n);
            if (!first) first = &o;
            else if (o.lc_n != first->lc_n) uniform = false;
            else for (int t = 0; t < o.lc_n && uniform; t++) uniform = p.lc_ser[o.lc_off + t] == p.lc_ser[first->lc_off + t];
        }
        /* clusters with a linear combination at this position: coefficients term-major over exactly those clusters */
        std::vector<int> lc_rank(ncl, -1);
        int nlc = 0;
        for (int c = 0; c < ncl; c++)
            if (pos < (int)cl_ops[c].size() && p.ops[cl_ops[c][pos]].op == W_LINCOMB) lc_rank[c] = nlc++;
        const size_t cbase = w->lc_coef.size(), sbase = w->lc_ser.size();
        if (longest > 0) {
            w->lc_coef.resize(cbase + (size_t)longest * nlc, 0.0);
            w->lc_ser.resize(sbase + (uniform ? (size_t)longest : (size_t)longest * ncl), 0);
        }
        for (int c = 0; c < ncl; c++) {
            C2Op &d = w->ops[(size_t)pos * ncl + c];
            d.op = W_NOP;
            const int nops_c = (int)cl_ops[c].size();
            if (pos >= nops_c + (int)cl_sol[c].size()) continue;
            if (pos >= nops_c) { /* solcoeff! */
                const int comp = cl_sol[c][pos - nops_c];
                int ng;
                d.op = W_SOLCOEFF; d.res = pid[p.xser[comp]]; d.a = loc(p.dxser[comp], &ng); d.b = -1; d.aux = -1; d.para = d.parb = -1;
                continue;
            }
            const WOp &o = p.ops[cl_ops[c][pos]];
            int na = 0, nb = 0, nx = 0;
            d.op = o.op; d.res = loc(o.res, &nx); d.a = loc(o.a, &na); d.b = loc(o.b, &nb); d.aux = o.aux >= 0 ? loc(o.aux, &nx) : -1;
            d.flags = (na ? 1 : 0) | (nb ? 2 : 0);
            d.para = o.para; d.parb = o.parb; d.imma = o.imma; d.immb = o.immb; d.imm = o.imm;
            if ((na || nb) && !(o.op == TB_OP_MUL && o.a >= 0 && o.b >= 0)) bad = "internal: negated operand outside a product";
            if (cauchy_op(o, 0) && d.a < -1) bad = "internal: recurrence operand is not stored";
            if (cauchy_op(o, 1) && d.b < -1) bad = "internal: recurrence operand is not stored";
            if (o.op == W_LINCOMB) {
                d.lc_n = o.lc_n; d.lc_off = (int)(cbase + lc_rank[c]); d.lc_stride = nlc;
                d.lc_soff = (int)(uniform ? sbase : sbase + c); d.lc_sstride = uniform ? 1 : ncl;
                d.pad = 0;
                for (int t = 0; t < o.lc_n; t++) {
                    int ng;
                    const int l = loc(p.lc_ser[o.lc_off + t], &ng);
                    if (l < 0 || ng) bad = "a linear-combination term is not a stored series"; /* terms are state jets in practice */
                    w->lc_coef[cbase + (size_t)t * nlc + lc_rank[c]] = p.lc_coef[o.lc_off + t];
                    w->lc_ser[uniform ? sbase + t : sbase + (size_t)t * ncl + c] = l;
                }
                {   /* equally spaced stored series (or a single term)? then the kernel walks a pointer, no id table */
                    auto sid_at = [&](int t) { return w->lc_ser[uniform ? sbase + t : sbase + (size_t)t * ncl + c]; };
                    const int s0 = sid_at(0), ds = o.lc_n >= 2 ? sid_at(1) - s0 : 0;
                    bool affine = true;
                    for (int t = 2; t < o.lc_n && affine; t++) affine = sid_at(t) == s0 + t * ds;
                    if (affine) { d.lc_sstride = 0; d.lc_soff = s0; d.pad = ds; }
                }
            }
        }
    }
    if (!bad.empty()) { why = bad; delete w; return nullptr; }
    /* GROUP PROGRAMS.  The clusters of a group run the same entries; with the role-major numbering their stored
     * operands differ by the cluster's rank in the group (or not at all), their linear-combination columns by the
     * rank, and their literal by a per-cluster table.  Then ONE copy of the entries serves the whole group: every
     * lane of a warp reads the same 80 bytes (a broadcast that stays in L1) instead of its own copy from L2 --
     * the per-cluster copies are 182 KB for many-spin 100 and their fetch latency was half of the step time.    */
    {
        bool ok = grp_count.size() <= 8;
        std::vector<C2Op> gops;
        std::vector<double> gnum;
        std::vector<int> goff, glen;
        for (size_t g = 0; g < grp_count.size() && ok; g++) {
            const int c0 = grp_first[g], cnt = grp_count[g], len = w->cl_len[c0];
            goff.push_back((int)gops.size());
            glen.push_back(len);
            for (int j = 0; j < len && ok; j++) {
                C2Op u = w->ops[(size_t)j * ncl + c0];
                int dflags = 0;
                const int num_off = (int)gnum.size();
                const bool sa = u.a == -1 && u.para < 0, sb = u.b == -1 && u.parb < 0;
                for (int r = 0; r < cnt && ok; r++) {
                    const C2Op &v = w->ops[(size_t)j * ncl + c0 + r];
                    auto delta = [&](int lu, int lv, int bit) {
                        if (lu < 0) { if (lv != lu) ok = false; return; }
                        const int dlt = lv - lu;
                        if (r == 1) { if (dlt == 1) dflags |= bit; else if (dlt != 0) ok = false; }
                        else if (dlt != ((dflags & bit) ? r : 0)) ok = false;
                    };
                    if (getenv("TB_WIDE_DEBUG") && r == 1) fprintf(stderr, "g%zu j%d op %d/%d fl %d/%d par %d/%d %d/%d n %d/%d ls %d/%d ss %d/%d pad %d/%d res %d/%d a %d/%d b %d/%d aux %d/%d off %d/%d soff %d/%d\n", g, j, u.op, v.op, u.flags, v.flags, u.para, v.para, u.parb, v.parb, u.lc_n, v.lc_n, u.lc_stride, v.lc_stride, u.lc_sstride, v.lc_sstride, u.pad, v.pad, u.res, v.res, u.a, v.a, u.b, v.b, u.aux, v.aux, u.lc_off, v.lc_off, u.lc_soff, v.lc_soff);
                    if (v.op != u.op || v.flags != u.flags || v.para != u.para || v.parb != u.parb || (u.op == TB_OP_POW && v.imm != u.imm) || v.lc_n != u.lc_n ||
                        v.lc_stride != u.lc_stride || v.lc_sstride != u.lc_sstride || v.pad != u.pad) { ok = false; break; }
                    delta(u.res, v.res, 16); delta(u.a, v.a, 32); delta(u.b, v.b, 64); delta(u.aux, v.aux, 128);
                    if (u.op == W_LINCOMB) {
                        if (v.lc_off != u.lc_off + r) ok = false;
                        if (u.lc_sstride == 0) delta(u.lc_soff, v.lc_soff, 256);          /* first operand series: fixed or by rank */
                        else if (u.lc_sstride == 1) { if (v.lc_soff != u.lc_soff) ok = false; } /* shared id table */
                        else ok = false;                                                  /* per-cluster id tables: keep per-cluster entries */
                    }
                    if (!ok && getenv("TB_WIDE_DEBUG")) fprintf(stderr, "FAIL g%zu j%d r%d: res %d/%d a %d/%d b %d/%d aux %d/%d dflags %d off %d/%d soff %d/%d\n", g, j, r, u.res, v.res, u.a, v.a, u.b, v.b, u.aux, v.aux, dflags, u.lc_off, v.lc_off, u.lc_soff, v.lc_soff);
                    gnum.push_back(sa ? v.imma : (sb ? v.immb : 0.0));
                }
                if (cnt == 1) dflags = 0;
                u.flags |= dflags;
                u.lc_soff = u.lc_soff;
                u.pad = u.pad;
                /* literal of cluster r at gnum[num_off + r]; stash the offset in the (unused here) immb slot as bits */
                long long no = num_off;
                memcpy(&u.immb, &no, sizeof no);
                gops.push_back(u);
            }
        }
        if (ok && !getenv("TB_WIDE_NO_GROUPS")) {
            w->gops.swap(gops); w->gnum.swap(gnum);
            for (size_t g = 0; g < grp_count.size(); g++) { w->g_first.push_back(grp_first[g]); w->g_count.push_back(grp_count[g]); }
            w->g_off = goff; w->g_len = glen;
        }
    }
    const int threads = threads_pick;
    w->threads = threads;
    w->smem = ((size_t)(p.order + 1) * NS + (size_t)w->ntr * threads + p.dof + 16) * sizeof(double);
    const size_t smem_max = 227 * 1024, smem_base = w->smem;
    /* the coupling constants are loop-invariant (every order of every step of every trajectory reads all of them):
     * they are copied to shared memory once per CTA when they fit beside the jets.  From global memory the dependent
     * add chain waits on L2 (measured on many-spin 100: 9.1e5 steps/s with 2 CTAs per SM and the table in L1/L2).  */
    w->lc_smem = !w->lc_coef.empty() && w->smem + w->lc_coef.size() * sizeof(double) + 1024 <= smem_max;
    /* ncu (profiles/r02_ncu_wide2_summary.txt): with the table in L1/L2 the multiplies of the coupling sum wait ~300
     * cycles per batch of loads, 25 batches per order; TB_WIDE_LC_GLOBAL=1 keeps it in global memory (2 CTAs per SM) */
    if (!getenv("TB_WIDE_LC_SMEM")) w->lc_smem = false; /* measured (many-spin 100, 2^14 members): 1.17e6 steps/s with the table in
                                                          * shared memory and 1 CTA per SM, 1.75e6 from L1/L2 with 2 CTAs per SM */
    if (w->lc_smem) w->smem += w->lc_coef.size() * sizeof(double);
    /* group programs and their literals live in shared memory as well (a few KB) */
    w->g_smem = !w->gops.empty() && w->smem + w->gops.size() * sizeof(C2Op) + w->gnum.size() * sizeof(double) + 1024 <= smem_max;
    if (w->g_smem) w->smem += w->gops.size() * sizeof(C2Op) + w->gnum.size() * sizeof(double);
    if (w->smem > smem_max) { why = "stored series do not fit shared memory (" + std::to_string(w->smem) + " B)"; delete w; return nullptr; }
    w->ctas = (int)std::max<size_t>(1, std::min<size_t>(smem_max / (w->smem + 1024), 2048 / threads));
    /* NVRTC specialisation of the group programs (TB_WIDE_INTERP=1 keeps the interpreter) */
    if (!w->gops.empty() && w->gops.size() <= 512 && !getenv("TB_WIDE_INTERP")) {
        const size_t lcb = w->lc_coef.size() * sizeof(double), gnb = w->gnum.size() * sizeof(double);
        auto ctas_for = [&](size_t bytes) { return (int)std::max<size_t>(1, std::min<size_t>(smem_max / (bytes + 1024), 2048 / threads)); };
        /* coupling constants in shared memory only when that does not cost a resident CTA: measured on many-spin 100
         * (2^14 members), 2 CTAs per SM with the constants through L1/L2 (the head of each sum in registers) 5.5e6
         * steps/s, 1 CTA per SM with all of them in shared memory 3.2e6 (one warp per scheduler: every latency is
         * exposed).  TB_WIDE_LC_SMEM=1 / TB_WIDE_LC_GLOBAL=1 force either placement.                                */
        w->gen_lc_smem = lcb > 0 && smem_base + lcb + gnb + 1024 <= smem_max &&
                         (ctas_for(smem_base + lcb + gnb) == ctas_for(smem_base + gnb) || getenv("TB_WIDE_LC_SMEM")) && !getenv("TB_WIDE_LC_GLOBAL");
        w->gen_smem = smem_base + (w->gen_lc_smem ? lcb : 0) + gnb;
        w->gen_ctas = ctas_for(w->gen_smem);
        if (const char *e = getenv("TB_WIDE_CTAS")) w->gen_ctas = std::max(1, std::min(w->gen_ctas, atoi(e)));
        int batch = w->gen_lc_smem ? 6 : 16;
        if (const char *e = getenv("TB_WIDE_LCB")) batch = std::max(1, std::min(32, atoi(e)));
        /* doubles of coupling constants per thread held in registers for the whole launch (measured on many-spin 100:
         * 0 -> 4.28e6, 24 -> 4.76e6, 44 -> 5.48e6, 56 -> 5.58e6, 64 -> 5.64e6 steps/s; 234 registers at 64) */
        int lcreg = w->gen_lc_smem ? 0 : 60;
        if (const char *e = getenv("TB_WIDE_LCREG")) lcreg = std::max(0, std::min(100, atoi(e)));
        w->gen_source = wide2_generate(w, w->gen_lc_smem, w->gen_ctas, batch, lcreg);
        if (tb::compile_cubin(w->gen_source, w->mode, !getenv("TB_WIDE_NO_CACHE"), w->gen_cubin, w->gen_log) != TB_OK) {
            if (getenv("TB_WIDE_DEBUG")) fprintf(stderr, "wide2: NVRTC specialisation failed, interpreter kept:\n%s\n", w->gen_log.c_str());
            w->gen_cubin.clear();
        } else {
            const size_t u = w->gen_log.find("Used ");
            if (u != std::string::npos) sscanf(w->gen_log.c_str() + u, "Used %d registers", &w->gen_regs);
            const size_t f = w->gen_log.find(" bytes stack frame");
            if (f != std::string::npos) {
                size_t b = f;
                while (b > 0 && isdigit((unsigned char)w->gen_log[b - 1])) b--;
                w->gen_local = atoi(w->gen_log.c_str() + b);
            }
        }
    }
    return w;
}

/* ---- NVRTC specialisation ------------------------------------------------------------------------------------
 * ncu of the interpreter (profiles/r02_ncu_wide2_summary.txt): 84 % of the executed instructions are dispatch and
 * operand arithmetic, 16 % FP64.  With the group programs emitted as straight-line exec2 calls whose entries are
 * compile-time constants the switch, the flag tests and the address arithmetic fold away, the coupling sum is
 * unrolled with its length known, and NS / P / the thread count are literals.  Same device text (tb_wide2_dev.cuh),
 * same operation order: bit-identical to the interpreter and to the oracle in parity mode.                       */
static std::string wide2_generate(const Wide2 *w, bool lc_smem, int ctas, int lcb, int lcreg) {
    std::string s;
    char buf[1024];
    auto line = [&](const char *fmt, auto... args) { snprintf(buf, sizeof buf, fmt, args...); s += buf; s += '\n'; };
    line("/* generated by libtaylorb200 (tb_wide2.cu, wide2_generate): cluster kernel, order %d, %d stored series, %d clusters */",
         w->order, w->NS, w->ncl);
    line("#define TB_W2_GENERATED %d", 1);
    line("#define TB_W2_NS %d", w->NS);
    line("#define TB_W2_P %d", w->order);
    line("#define TB_W2_THREADS %d", w->threads);
    line("#define TB_W2_LCB %d", lcb);
    line("#define TB_W2_LC_SMEM %d", lc_smem ? 1 : 0);
    /* register cache of coupling constants: short sums first (their single load would sit on the critical path of
     * every order), then the head of the long ones, `lcreg` doubles per thread in all */
    struct Cached { size_t g; int j, n, off; };
    std::vector<Cached> cached;
    {
        int left = lcreg;
        for (int pass = 0; pass < 2 && left > 0; pass++)
            for (size_t g = 0; g < w->g_count.size() && left > 0; g++)
                for (int j = 0; j < w->g_len[g] && left > 0; j++) {
                    const C2Op &o = w->gops[w->g_off[g] + j];
                    if (o.op != W_LINCOMB || o.lc_sstride != 0 || o.lc_n <= 0) continue;
                    if ((pass == 0) != (o.lc_n <= 4)) continue;
                    const int n = std::min(o.lc_n, left);
                    cached.push_back({g, j, n, lcreg - left});
                    left -= n;
                }
        lcreg -= left;
    }
    auto find_cached = [&](size_t g, int j) -> const Cached * { for (auto &c : cached) if (c.g == g && c.j == j) return &c; return nullptr; };
    line("#define TB_W2_LCREG %d", lcreg);
    s += "enum {";
    const std::pair<const char *, int> ops[] = {
        {"TB_OP_NOP", TB_OP_NOP}, {"TB_OP_IDENTITY", TB_OP_IDENTITY}, {"TB_OP_NEG", TB_OP_NEG}, {"TB_OP_ADD", TB_OP_ADD},
        {"TB_OP_SUB", TB_OP_SUB}, {"TB_OP_MUL", TB_OP_MUL}, {"TB_OP_DIV", TB_OP_DIV}, {"TB_OP_SQR", TB_OP_SQR},
        {"TB_OP_SQRT", TB_OP_SQRT}, {"TB_OP_POW", TB_OP_POW}, {"TB_OP_EXP", TB_OP_EXP}, {"TB_OP_LOG", TB_OP_LOG},
        {"TB_OP_SINCOS", TB_OP_SINCOS}, {"TB_OP_ZERO", TB_OP_ZERO}, {"TB_OP_ONE", TB_OP_ONE}, {"W_NOP", W_NOP},
        {"W_LINCOMB", W_LINCOMB}, {"W_SOLCOEFF", W_SOLCOEFF}, {"W_MUL2", W_MUL2}, {"TB_ST_MAXSTEPS", TB_ST_MAXSTEPS},
        {"TB_ST_ZERO_STEP", TB_ST_ZERO_STEP}, {"TB_ST_NONFINITE", TB_ST_NONFINITE}};
    for (auto &kv : ops) { snprintf(buf, sizeof buf, " %s = %d,", kv.first, kv.second); s += buf; }
    s += " TB_W2_ENUM_END };\n";
    s += tb::embedded_math_h();
    s += tb::embedded_wide2_dev();
    line("%s", "template <bool FAST>");
    line("%s", "__device__ __forceinline__ void tb_w2_groups(const W2Args &a, double *__restrict__ W, double *__restrict__ tr, const double *par,");
    line("%s", "                                             const double *__restrict__ lcc, const double *__restrict__ gnum, const int k, const int tid,");
    line("%s", "                                             const double *lcreg) {");
    for (size_t g = 0; g < w->g_count.size(); g++) {
        line("  /* group %zu: %d clusters x %d entries */", g, w->g_count[g], w->g_len[g]);
        if (w->g_count[g] <= w->threads) line("  if (tid < %d) { const int r = tid;", w->g_count[g]);
        else line("  for (int r = tid; r < %d; r += TB_W2_THREADS) {", w->g_count[g]);
        /* fuse two Cauchy products that share their second operand (many-spin: z_i * S_i and -y_i * S_i) into one
         * W_MUL2 entry at the place of the first: the later one may move up when no entry in between writes what it
         * reads or touches what it writes.  Locations compare as (value, shifts-with-rank): equal values, or
         * persistent ids with different shift behaviour, count as possibly the same series. */
        const int glen = w->g_len[g];
        std::vector<int> partner(glen, -1);
        std::vector<char> skip(glen, 0);
        if (!getenv("TB_WIDE_NO_MUL2")) {
            auto E = [&](int j) -> const C2Op & { return w->gops[w->g_off[g] + j]; };
            auto cauchy = [&](const C2Op &o) { return o.op == TB_OP_MUL && o.a >= 0 && o.b >= 0; };
            auto same = [&](int va, bool sa, int vb, bool sb) {
                if (va == -1 || vb == -1) return false;          /* numbers */
                if (va < 0 || vb < 0) return va == vb;            /* transient slots: per thread, by number */
                return va == vb || sa != sb;                      /* persistent ids */
            };
            for (int p0 = 0; p0 < glen; p0++) {
                if (skip[p0] || partner[p0] >= 0 || !cauchy(E(p0))) continue;
                const C2Op &u = E(p0);
                for (int q = p0 + 1; q < glen; q++) {
                    const C2Op &v = E(q);
                    if (skip[q] || !cauchy(v) || v.b != u.b || ((v.flags & 64) != (u.flags & 64)) || ((v.flags & 2) != (u.flags & 2))) continue;
                    bool ok = !same(v.a, v.flags & 32, u.res, u.flags & 16) && !same(v.res, v.flags & 16, u.res, u.flags & 16);
                    for (int m = p0 + 1; m < q && ok; m++) {
                        const C2Op &x = E(m);
                        if (skip[m]) continue;
                        const bool xs16 = x.flags & 16, xs128 = x.flags & 128;
                        /* writes of x: res (and aux for sincos-like entries) against what v reads (a, b) and writes (res) */
                        const int wr[2] = {x.res, x.aux};
                        const bool ws[2] = {xs16, xs128};
                        for (int t = 0; t < 2 && ok; t++) {
                            if (t == 1 && x.op != TB_OP_SINCOS) break;
                            if (same(wr[t], ws[t], v.a, v.flags & 32) || same(wr[t], ws[t], v.b, v.flags & 64) ||
                                same(wr[t], ws[t], v.res, v.flags & 16)) ok = false;
                        }
                        if (x.op == W_SOLCOEFF) ok = ok && !same(x.res, xs16, v.a, v.flags & 32) && !same(x.res, xs16, v.b, v.flags & 64);
                        /* reads of x against v's result; a coupling sum reads persistent series by table */
                        if (same(x.a, x.flags & 32, v.res, v.flags & 16) || same(x.b, x.flags & 64, v.res, v.flags & 16)) ok = false;
                        if (x.op == W_LINCOMB && v.res >= 0) ok = false;
                    }
                    if (ok) { partner[p0] = q; skip[q] = 1; break; }
                }
            }
        }
        for (int j = 0; j < w->g_len[g]; j++) {
            if (skip[j]) continue;
            C2Op o = w->gops[w->g_off[g] + j];
            if (partner[j] >= 0) { /* second product: a2 -> aux, res2 -> lc_soff, its negation -> bit 2 */
                const C2Op &v = w->gops[w->g_off[g] + partner[j]];
                o.op = W_MUL2;
                o.aux = v.a; o.lc_soff = v.res;
                o.flags = (o.flags & ~(128 | 256 | 4)) | ((v.flags & 32) ? 128 : 0) | ((v.flags & 16) ? 256 : 0) | ((v.flags & 1) ? 4 : 0);
            }
            long long no;
            memcpy(&no, &o.immb, sizeof no);
            const bool lit = (o.a == -1 && o.para < 0) || (o.b == -1 && o.parb < 0);
            auto fld = [&](int v, int bit) { return std::to_string(v) + ((o.flags & bit) ? " + r" : ""); };
            std::string e = "    { C2Op o; o.op = " + std::to_string(o.op) + "; o.res = " + fld(o.res, 16) + "; o.a = " + fld(o.a, 32) +
                            "; o.b = " + fld(o.b, 64) + "; o.aux = " + fld(o.aux, 128) + "; o.flags = " + std::to_string(o.flags & 15) +
                            "; o.para = " + std::to_string(o.para) + "; o.parb = " + std::to_string(o.parb) + ";";
            s += e + "\n";
            snprintf(buf, sizeof buf, "      o.lc_off = %d + r; o.lc_n = %d; o.lc_stride = %d; o.lc_soff = %s; o.lc_sstride = %d; o.pad = %d; o.imm = %a;",
                     o.lc_off, o.lc_n, o.lc_stride, fld(o.lc_soff, 256).c_str(), o.lc_sstride, o.pad, o.imm);
            s += buf; s += '\n';
            if (lit) line("      o.imma = o.immb = gnum[%lld + r];", no);
            else line("      o.imma = o.immb = %s;", "0.0");
            if (const Cached *c = find_cached(g, j)) line("      exec2<FAST>(a, W, tr, par, lcc, %s, o, k, lcreg + %d, %d); }", lc_smem ? "true" : "false", c->off, c->n);
            else line("      exec2<FAST>(a, W, tr, par, lcc, %s, o, k); }", lc_smem ? "true" : "false");
        }
        line("  %s", "}");
    }
    line("%s", "}");
    line("%s", "__device__ __forceinline__ void tb_w2_lcreg_fill(double *lcreg, const double *__restrict__ lcc, const int tid) {");
    for (auto &c : cached) {
        const C2Op &o = w->gops[w->g_off[c.g] + c.j];
        line("  { const int r = tid < %d ? tid : 0; /* group %zu entry %d: %d of %d terms */", w->g_count[c.g], c.g, c.j, c.n, o.lc_n);
        line("    _Pragma(\"unroll\") for (int u = 0; u < %d; u++) lcreg[%d + u] = lcc[%d + r + u * %d]; }", c.n, c.off, o.lc_off, o.lc_stride);
    }
    line("%s", "}");
    line("extern \"C\" __global__ void __launch_bounds__(%d, %d) tb_w2_kernel(const W2Args a) { wide2_body<%s>(a); }", w->threads, ctas,
         w->mode == TB_MODE_FAST ? "true" : "false");
    return s;
}

void wide2_free(Wide2 *w) {
    if (!w) return;
    for (auto &kv : w->gen_loaded) tb::module_unload(kv.second.first);
    if (w->d_ops) { cudaFree(w->d_ops); cudaFree(w->d_lcc); cudaFree(w->d_lcs); cudaFree(w->d_xs); cudaFree(w->d_len); cudaFree(w->d_gops); cudaFree(w->d_gnum); }
    delete w;
}

void wide2_info(const Wide2 *w, int *nclusters, int *npersist, int *ntransient, int *threads, int *smem_bytes, int *ctas_per_sm, int *ngroups) {
    if (!w) return;
    if (nclusters) *nclusters = w->ncl;
    if (npersist) *npersist = w->NS;
    if (ntransient) *ntransient = w->ntr;
    if (threads) *threads = w->threads;
    const bool gen = !w->gen_cubin.empty();
    if (smem_bytes) *smem_bytes = (int)(gen ? w->gen_smem : w->smem);
    if (ctas_per_sm) *ctas_per_sm = gen ? w->gen_ctas : w->ctas;
    if (ngroups) *ngroups = w->gops.empty() ? 0 : (int)w->g_count.size();
}

void wide2_gen_info(const Wide2 *w, int *generated, int *regs, int *local_bytes, int *lc_in_smem) {
    const bool gen = w && !w->gen_cubin.empty();
    if (generated) *generated = gen ? 1 : 0;
    if (regs) *regs = gen ? w->gen_regs : 0;
    if (local_bytes) *local_bytes = gen ? w->gen_local : 0;
    if (lc_in_smem) *lc_in_smem = gen ? (w->gen_lc_smem ? 1 : 0) : (w && w->lc_smem ? 1 : 0);
}
const char *wide2_gen_source(const Wide2 *w) { return w ? w->gen_source.c_str() : nullptr; }
const char *wide2_gen_log(const Wide2 *w) { return w ? w->gen_log.c_str() : nullptr; }

int wide2_integrate(Wide2 *w, long long N, const double *q0, const double *params, double t0, double tmax,
                    const double *trange, long long nt, const tb_opts_t *o, const tb_out_t *out) {
    int dev = 0, nsm = 148;
    TB_CUDA2(cudaGetDevice(&dev));
    TB_CUDA2(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    const int dof = w->dof;
    const bool fast = w->mode == TB_MODE_FAST;
    if (!w->d_ops || w->device != dev) {
        if (w->d_ops) { cudaFree(w->d_ops); cudaFree(w->d_lcc); cudaFree(w->d_lcs); cudaFree(w->d_xs); cudaFree(w->d_len); cudaFree(w->d_gops); cudaFree(w->d_gnum); w->d_ops = nullptr; }
        w->device = dev;
        auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
            cudaError_t e = cudaMalloc(dst, bytes ? bytes : 8);
            if (e != cudaSuccess) return e;
            return bytes ? cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
        };
        TB_CUDA2(up((void **)&w->d_ops, w->ops.data(), sizeof(C2Op) * w->ops.size()));
        TB_CUDA2(up((void **)&w->d_lcc, w->lc_coef.data(), sizeof(double) * w->lc_coef.size()));
        TB_CUDA2(up((void **)&w->d_lcs, w->lc_ser.data(), sizeof(int) * w->lc_ser.size()));
        TB_CUDA2(up((void **)&w->d_xs, w->xsid.data(), sizeof(int) * dof));
        TB_CUDA2(up((void **)&w->d_len, w->cl_len.data(), sizeof(int) * w->cl_len.size()));
        TB_CUDA2(up((void **)&w->d_gops, w->gops.data(), sizeof(C2Op) * w->gops.size()));
        TB_CUDA2(up((void **)&w->d_gnum, w->gnum.data(), sizeof(double) * w->gnum.size()));
        if (fast) TB_CUDA2(cudaFuncSetAttribute(wide2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w->smem));
        else TB_CUDA2(cudaFuncSetAttribute(wide2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w->smem));
    }
    const bool gen = !w->gen_cubin.empty();
    void *gen_fn = nullptr;
    if (gen) { /* the specialised kernel: one module per device */
        std::lock_guard<std::mutex> lk(w->gen_mu);
        auto it = w->gen_loaded.find(dev);
        if (it == w->gen_loaded.end()) {
            void *mod = nullptr, *fn = nullptr;
            const int lrc = tb::module_load(w->gen_cubin, "tb_w2_kernel", w->gen_smem, &mod, &fn);
            if (lrc != TB_OK) return lrc;
            it = w->gen_loaded.emplace(dev, std::make_pair(mod, fn)).first;
        }
        gen_fn = it->second.second;
    }
    int bps = 1;
    if (gen) bps = w->gen_ctas;
    else if (fast) TB_CUDA2(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide2_kernel<true>, w->threads, w->smem));
    else TB_CUDA2(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide2_kernel<false>, w->threads, w->smem));
    bps = std::max(1, bps);
    const long long blocks = std::min<long long>((long long)nsm * bps, N);
    W2Args a{};
    a.N = N; a.dof = dof; a.npar = w->npar; a.P = w->order; a.NS = w->NS; a.ncl = w->ncl; a.maxlen = w->maxlen; a.ntr = w->ntr; a.tsid = w->tsid;
    a.ops = w->d_ops; a.lc_coef = w->d_lcc; a.lc_ser = w->d_lcs; a.xsid = w->d_xs; a.cl_len = w->d_len; a.lc_smem = w->lc_smem ? 1 : 0; a.lc_count = (int)w->lc_coef.size();
    a.ngroups = w->gops.empty() ? 0 : (int)w->g_count.size(); a.gops = w->d_gops; a.gnum = w->d_gnum;
    a.g_smem = w->g_smem ? 1 : 0; a.ngops = (int)w->gops.size(); a.ngnum = (int)w->gnum.size();
    if (gen) { a.lc_smem = w->gen_lc_smem ? 1 : 0; a.g_smem = w->gnum.empty() ? 0 : 1; a.ngops = 0; }
    for (int g = 0; g < a.ngroups; g++) { a.g_first[g] = w->g_first[g]; a.g_count[g] = w->g_count[g]; a.g_off[g] = w->g_off[g]; a.g_len[g] = w->g_len[g]; }
    a.nt = trange ? nt : 0; a.t0 = t0; a.tmax = tmax; a.abstol = o->abstol; a.reltol = o->reltol; a.minstep = o->minstepsize;
    a.maxstep = o->maxstepsize; a.maxsteps = o->maxsteps; a.params_per_traj = o->params_per_traj;
    const size_t nq = (size_t)dof * N, npar = (size_t)(w->npar > 0 ? w->npar : 1) * (o->params_per_traj ? N : 1);
    double *dq = nullptr, *dp = nullptr, *dtr = nullptr, *dxf = nullptr, *dtf = nullptr, *dxv = nullptr;
    long long *dns = nullptr; int *dst = nullptr; unsigned long long *dctr = nullptr;
    int rc = TB_OK;
    auto fail = [&](cudaError_t e, const char *what) { if (e != cudaSuccess && rc == TB_OK) { tb::set_error(std::string(what) + ": " + cudaGetErrorString(e)); rc = TB_ERR_CUDA; } };
    fail(cudaMalloc(&dq, 8 * nq), "cudaMalloc q0");
    fail(cudaMalloc(&dp, 8 * npar), "cudaMalloc params");
    fail(cudaMalloc(&dxf, 8 * nq), "cudaMalloc xfinal");
    fail(cudaMalloc(&dtf, 8 * N), "cudaMalloc tfinal");
    fail(cudaMalloc(&dns, 8 * N), "cudaMalloc nsteps");
    fail(cudaMalloc(&dst, 4 * N), "cudaMalloc status");
    fail(cudaMalloc(&dctr, 8), "cudaMalloc counter");
    if (trange) {
        fail(cudaMalloc(&dtr, 8 * nt), "cudaMalloc trange");
        if (out->xv_trange) fail(cudaMalloc(&dxv, 8 * nq * nt), "cudaMalloc xv_trange");
    }
    if (rc == TB_OK) {
        fail(cudaMemcpy(dq, q0, 8 * nq, cudaMemcpyHostToDevice), "copy q0");
        if (w->npar > 0 && params) fail(cudaMemcpy(dp, params, 8 * npar, cudaMemcpyHostToDevice), "copy params");
        if (trange) fail(cudaMemcpy(dtr, trange, 8 * nt, cudaMemcpyHostToDevice), "copy trange");
        fail(cudaMemset(dctr, 0, 8), "memset counter");
        a.q0 = dq; a.params = dp; a.trange = trange ? dtr : nullptr; a.xfinal = dxf; a.tfinal = dtf;
        a.nsteps = dns; a.status = dst; a.xv_trange = dxv; a.counter = dctr;
    }
    if (rc == TB_OK) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, 0);
        if (gen) { const int lrc = tb::launch_fn(gen_fn, (unsigned)blocks, (unsigned)w->threads, w->gen_smem, nullptr, &a); if (lrc != TB_OK) rc = lrc; }
        else if (fast) wide2_kernel<true><<<(unsigned)blocks, w->threads, w->smem>>>(a);
        else wide2_kernel<false><<<(unsigned)blocks, w->threads, w->smem>>>(a);
        cudaEventRecord(e1, 0);
        if (!gen) tb::count_launch(1);
        fail(cudaGetLastError(), "wide2_kernel launch");
        fail(cudaDeviceSynchronize(), "wide2_kernel");
        float ms = -1.f;
        cudaEventElapsedTime(&ms, e0, e1);
        tb::set_last_kernel_ms(ms);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    if (rc == TB_OK) {
        if (out->xfinal) fail(cudaMemcpy(out->xfinal, dxf, 8 * nq, cudaMemcpyDeviceToHost), "copy xfinal");
        if (out->tfinal) fail(cudaMemcpy(out->tfinal, dtf, 8 * N, cudaMemcpyDeviceToHost), "copy tfinal");
        if (out->nsteps) fail(cudaMemcpy(out->nsteps, dns, 8 * N, cudaMemcpyDeviceToHost), "copy nsteps");
        if (out->status) fail(cudaMemcpy(out->status, dst, 4 * N, cudaMemcpyDeviceToHost), "copy status");
        if (out->xv_trange && dxv) fail(cudaMemcpy(out->xv_trange, dxv, 8 * nq * nt, cudaMemcpyDeviceToHost), "copy xv_trange");
    }
    cudaFree(dq); cudaFree(dp); cudaFree(dtr); cudaFree(dxf); cudaFree(dtf); cudaFree(dxv);
    cudaFree(dns); cudaFree(dst); cudaFree(dctr);
    return rc;
}

} // namespace tbw
