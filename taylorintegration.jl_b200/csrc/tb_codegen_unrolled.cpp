/*
 * tb_codegen_unrolled.cpp -- the register-tiled code generator (kernel variant 2).
 *
 * Same lowering as tb_codegen.cpp (one TaylorSeries.<op>!(res, a, b, ord) per tape
 * entry, src/parse_eqs.jl:963-1135, run for ord = 0..order-1, parse_eqs.jl:259-264,
 * then solcoeff!, src/TaylorIntegration.jl:33-40), but the `ord` loop and every
 * recurrence sum are fully unrolled at code-generation time.  All coefficient
 * indices are therefore compile-time constants, and a stored series can live in
 * REGISTERS (a struct of arrays that ptxas scalarises) instead of shared memory.
 *
 * Why: a DFMA whose two operands come from shared memory is capped at 1/8 of the
 * FP64 pipe by the 128 B/clk crossbar (B200: 64 DFMA/clk/SM vs 16 doubles/clk),
 * and the rolled loops of variant 1 are latency-bound besides.  Placement is by
 * read count: the series read most often by Cauchy sums go to registers until
 * the budget is spent, the rest to shared memory.  In fast mode a state jet that
 * no recurrence reads and whose derivative jet is stored (x[k] = dx[k-1]/k) is
 * not stored at all.
 *
 * Parity mode emits the oracle's operation order with unfused multiply/add and
 * is bit-identical to variant 1; fast mode uses FMA, 4 partial accumulators per
 * sum and reciprocal multiplies.
 */
#include "tb_internal.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <set>

namespace tb {

namespace {

struct Em {
    std::string s;
    void line(const char *fmt, ...) {
        char buf[4096];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        s += buf;
        s += "\n";
    }
};

std::string dlit(double v) {
    unsigned long long u;
    memcpy(&u, &v, 8);
    char buf[96];
    if (v == (double)(long long)v && std::fabs(v) < 1e15) snprintf(buf, sizeof buf, "%.1f", v);
    else snprintf(buf, sizeof buf, "__longlong_as_double(0x%016llxLL)", u);
    return buf;
}
std::string fmt(const char *f, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, f);
    vsnprintf(buf, sizeof buf, f, ap);
    va_end(ap);
    return buf;
}

bool u_is_num(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_PAR || k == TB_K_IMM; }
bool u_is_series(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_X || k == TB_K_DX || k == TB_K_TMP || k == TB_K_TIME; }

enum Place { PL_NONE = 0, PL_REG = 1, PL_SMEM = 2, PL_DERIVED = 3 };

struct SeriesInfo {
    int place = PL_NONE;
    int id = -1;     /* register array index or shared-memory slot */
    double reads = 0; /* Cauchy-sum reads per step (placement weight) */
    int len = 0;      /* coefficients kept */
};

struct UGen {
    const Ode &ode;
    const tb_compile_opts_t &co;
    Layout &L;
    const int P;
    const bool fast;
    std::map<int32_t, SeriesInfo> ser;
    std::string err;
    int nreg = 0, nsm = 0;
    int tmpc = 0;

    UGen(const Ode &o, const tb_compile_opts_t &c, Layout &l) : ode(o), co(c), L(l), P(c.order), fast(c.mode == TB_MODE_FAST) {}

    /* ---- analysis ---------------------------------------------------------- */
    void need(int32_t o, double w) {
        if (!u_is_series(o) || TB_OPERAND_KIND(o) == TB_K_TIME) return;
        ser[o].reads += w;
    }
    void analyse(int reg_budget) {
        const tb_ode_desc_t &d = ode.desc;
        const double w = 0.5 * P * P;
        for (int j = 0; j < d.nops; j++) {
            const tb_op_t &op = ode.ops[j];
            switch (op.op) {
                case TB_OP_MUL: if (u_is_series(op.a) && u_is_series(op.b)) { need(op.a, w); need(op.b, w); } break;
                case TB_OP_SQR: need(op.a, w); break;
                case TB_OP_DIV: if (u_is_series(op.b)) { need(op.b, w); need(op.res, w); } break;
                case TB_OP_SQRT: need(op.res, 0.5 * w); break;
                case TB_OP_POW:
                    if (op.imm == 2.0) need(op.a, w);
                    else if (op.imm == 0.5) need(op.res, 0.5 * w);
                    else if (op.imm != 0.0 && op.imm != 1.0) { need(op.a, w); need(op.res, w); }
                    break;
                case TB_OP_EXP: case TB_OP_LOG: need(op.a, w); need(op.res, w); break;
                case TB_OP_SINCOS: need(op.a, w); need(op.res, w); need(op.aux, w); break;
                default: break;
            }
        }
        /* state jets: stored unless derivable from a stored dx (fast mode only) */
        for (int c = 0; c < d.dof; c++) {
            int32_t x = TB_OPERAND(TB_K_X, c), dx = TB_OPERAND(TB_K_DX, c);
            bool cauchy = ser.count(x) && ser[x].reads > 0;
            bool dx_stored = ser.count(dx) && ser[dx].reads > 0;
            if (!cauchy && fast && dx_stored) { ser[x].place = PL_DERIVED; ser[x].id = c; }
            else ser[x].reads += 0.0; /* creates the entry: stored */
        }
        /* placement by read count */
        std::vector<std::pair<double, int32_t>> order;
        for (auto &kv : ser)
            if (kv.second.place != PL_DERIVED) order.push_back({-kv.second.reads, kv.first});
        std::sort(order.begin(), order.end());
        int used = 0;
        for (auto &pr : order) {
            SeriesInfo &si = ser[pr.second];
            si.len = P + 1;
            if (si.reads > 0 && used + si.len <= reg_budget) { si.place = PL_REG; si.id = nreg++; used += si.len; }
            else { si.place = PL_SMEM; si.id = nsm++; }
        }
    }

    /* ---- operand access ------------------------------------------------------ */
    std::string sref(int32_t o, int idx) {
        auto it = ser.find(o);
        if (it == ser.end() || it->second.place == PL_NONE || it->second.place == PL_DERIVED) {
            err = "internal: series operand not stored";
            return "0.0";
        }
        if (it->second.place == PL_REG) return fmt("J.r%d[%d]", it->second.id, idx);
        return fmt("S(%d, %d)", it->second.id, idx);
    }
    static std::string vname(int32_t o, int k) {
        int kind = TB_OPERAND_KIND(o), i = TB_OPERAND_INDEX(o);
        return fmt("%s%d_%d", kind == TB_K_DX ? "vD" : (kind == TB_K_TMP ? "vT" : "vX"), i, k);
    }
    std::string num(int32_t o, const tb_op_t &op) const {
        if (TB_OPERAND_KIND(o) == TB_K_PAR) return fmt("par[%d]", TB_OPERAND_INDEX(o));
        return dlit(op.imm);
    }
    /* value of series operand o at static index idx while generating order k */
    std::string rd(int32_t o, int idx, int k, const std::set<int32_t> &defined) {
        int kind = TB_OPERAND_KIND(o);
        if (kind == TB_K_TIME) return idx == 0 ? "t0" : (idx == 1 ? "1.0" : "0.0");
        if (idx == k && (kind == TB_K_X || defined.count(o))) return vname(o, k);
        return sref(o, idx);
    }
    bool is_zero_time(int32_t o, int idx) const { return TB_OPERAND_KIND(o) == TB_K_TIME && idx >= 2; }

    /* sum of products.  Parity: c = t0; c = c + t1; ... with unfused ops, in order.
     * Fast: up to 4 partial accumulators, FMA.  Each term = (x, y) meaning x*y;
     * `neg` subtracts the products instead (sincos cosine chain).                 */
    std::string sum(Em &e, const std::vector<std::pair<std::string, std::string>> &t, bool neg_chain = false) {
        std::string acc = fmt("a%d", tmpc++);
        if (t.empty()) { e.line("  const double %s = 0.0;", acc.c_str()); return acc; }
        if (!fast) {
            if (neg_chain) e.line("  double %s = M_(N_(%s), %s);", acc.c_str(), t[0].first.c_str(), t[0].second.c_str());
            else e.line("  double %s = M_(%s, %s);", acc.c_str(), t[0].first.c_str(), t[0].second.c_str());
            for (size_t i = 1; i < t.size(); i++)
                e.line("  %s = %s(%s, M_(%s, %s));", acc.c_str(), neg_chain ? "B_" : "A_", acc.c_str(), t[i].first.c_str(), t[i].second.c_str());
            return acc;
        }
        const int nacc = (int)std::min<size_t>(4, t.size());
        for (int j = 0; j < nacc; j++)
            e.line("  double %s_%d = %s * %s;", acc.c_str(), j, t[j].first.c_str(), t[j].second.c_str());
        for (size_t i = nacc; i < t.size(); i++)
            e.line("  %s_%d = fma(%s, %s, %s_%d);", acc.c_str(), (int)(i % nacc), t[i].first.c_str(), t[i].second.c_str(), acc.c_str(), (int)(i % nacc));
        std::string comb;
        if (nacc == 1) comb = fmt("%s_0", acc.c_str());
        else if (nacc == 2) comb = fmt("%s_0 + %s_1", acc.c_str(), acc.c_str());
        else if (nacc == 3) comb = fmt("(%s_0 + %s_1) + %s_2", acc.c_str(), acc.c_str(), acc.c_str());
        else comb = fmt("(%s_0 + %s_1) + (%s_2 + %s_3)", acc.c_str(), acc.c_str(), acc.c_str(), acc.c_str());
        e.line("  const double %s = %s(%s);", acc.c_str(), neg_chain ? "-" : "", comb.c_str());
        return acc;
    }

    typedef std::vector<std::pair<std::string, std::string>> Terms;
    void term(Terms &t, const std::string &x, const std::string &y) {
        if (fast && (x == "0.0" || y == "0.0")) return; /* structural zeros of the time series */
        t.push_back({x, y});
    }

    /* one tape entry at static order k */
    void emit_op(Em &e, int opi, const tb_op_t &op, int k, std::set<int32_t> &defined) {
        const bool sa = u_is_num(op.a), sb = u_is_num(op.b);
        const bool k0 = k == 0;
        const std::string R = vname(op.res, k);
        auto A = [&](int idx) { return sa ? num(op.a, op) : rd(op.a, idx, k, defined); };
        auto B = [&](int idx) { return sb ? num(op.b, op) : rd(op.b, idx, k, defined); };
        auto C = [&](int idx) { return rd(op.res, idx, k, defined); };
        const char *r = R.c_str();
        int opc = op.op;
        const double pw = op.imm;
        if (opc == TB_OP_POW) {
            if (pw == 0.0) opc = TB_OP_ONE;
            else if (pw == 1.0) opc = TB_OP_IDENTITY;
            else if (pw == 2.0) opc = TB_OP_SQR;
            else if (pw == 0.5) opc = TB_OP_SQRT;
        }
        switch (opc) {
            case TB_OP_IDENTITY:
                if (sa) e.line("  const double %s = %s;", r, k0 ? A(0).c_str() : "0.0");
                else e.line("  const double %s = %s;", r, A(k).c_str());
                break;
            case TB_OP_NEG: e.line("  const double %s = N_(%s);", r, A(k).c_str()); break;
            case TB_OP_ZERO: e.line("  const double %s = 0.0;", r); break;
            case TB_OP_ONE: e.line("  const double %s = %s;", r, k0 ? "1.0" : "0.0"); break;
            case TB_OP_ADD:
            case TB_OP_SUB: {
                const char *F = opc == TB_OP_ADD ? "A_" : "B_";
                if (sa && sb) { if (k0) e.line("  const double %s = %s(%s, %s);", r, F, A(0).c_str(), B(0).c_str()); else e.line("  const double %s = 0.0;", r); }
                else if (sa) {
                    if (k0) e.line("  const double %s = %s(%s, %s);", r, F, A(0).c_str(), B(0).c_str());
                    else if (opc == TB_OP_ADD) e.line("  const double %s = %s;", r, B(k).c_str());
                    else e.line("  const double %s = N_(%s);", r, B(k).c_str());
                } else if (sb) {
                    if (k0) e.line("  const double %s = %s(%s, %s);", r, F, A(0).c_str(), B(0).c_str());
                    else e.line("  const double %s = %s;", r, A(k).c_str());
                } else e.line("  const double %s = %s(%s, %s);", r, F, A(k).c_str(), B(k).c_str());
                break;
            }
            case TB_OP_MUL:
                if (sa && sb) { if (k0) e.line("  const double %s = M_(%s, %s);", r, A(0).c_str(), B(0).c_str()); else e.line("  const double %s = 0.0;", r); }
                else if (sa || sb) e.line("  const double %s = M_(%s, %s);", r, A(k).c_str(), B(k).c_str());
                else {
                    Terms t;
                    for (int i = 0; i <= k; i++) term(t, A(i), B(k - i));
                    std::string s = sum(e, t);
                    e.line("  const double %s = %s;", r, s.c_str());
                }
                break;
            case TB_OP_SQR:
                if (k0) e.line("  const double %s = M_(%s, %s);", r, A(0).c_str(), A(0).c_str());
                else {
                    const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                    Terms t;
                    for (int i = 0; i <= kend; i++) term(t, A(i), A(k - i));
                    std::string s = sum(e, t);
                    if (kodd) e.line("  const double %s = M_(2.0, %s);", r, s.c_str());
                    else if (fast) e.line("  const double %s = fma(%s, %s, 2.0 * %s);", r, A(k / 2).c_str(), A(k / 2).c_str(), s.c_str());
                    else e.line("  const double %s = A_(M_(2.0, %s), M_(%s, %s));", r, s.c_str(), A(k / 2).c_str(), A(k / 2).c_str());
                }
                break;
            case TB_OP_DIV:
                if (sa && sb) { if (k0) e.line("  const double %s = D_(%s, %s);", r, A(0).c_str(), B(0).c_str()); else e.line("  const double %s = 0.0;", r); }
                else if (sb) {
                    if (fast) { if (k0) e.line("  const double inv%d = 1.0 / %s;", opi, B(0).c_str()); e.line("  const double %s = %s * inv%d;", r, A(k).c_str(), opi); }
                    else e.line("  const double %s = D_(%s, %s);", r, A(k).c_str(), B(k).c_str());
                } else if (k0) {
                    if (fast) { e.line("  const double inv%d = 1.0 / %s;", opi, B(0).c_str()); e.line("  const double %s = %s * inv%d;", r, A(0).c_str(), opi); }
                    else {
                        e.line("  const double den%d = %s, inv%d = D_(1.0, den%d); const bool ok%d = tb_midrange(den%d);", opi, B(0).c_str(), opi, opi, opi, opi);
                        e.line("  const double %s = D_(%s, den%d);", r, A(0).c_str(), opi);
                    }
                } else {
                    Terms t;
                    for (int i = 0; i <= k - 1; i++) term(t, C(i), B(k - i));
                    std::string s = sum(e, t);
                    if (fast) {
                        if (sa) e.line("  const double %s = -%s * inv%d;", r, s.c_str(), opi);
                        else e.line("  const double %s = (%s - %s) * inv%d;", r, A(k).c_str(), s.c_str(), opi);
                    } else {
                        if (sa) e.line("  const double %s = tb_div_y(N_(%s), den%d, inv%d, ok%d);", r, s.c_str(), opi, opi, opi);
                        else e.line("  const double %s = tb_div_y(B_(%s, %s), den%d, inv%d, ok%d);", r, A(k).c_str(), s.c_str(), opi, opi, opi);
                    }
                }
                break;
            case TB_OP_SQRT:
                if (k0) {
                    e.line("  const double %s = Q_(%s);", r, A(0).c_str());
                    if (fast) e.line("  const double inv%d = 0.5 / %s;", opi, r);
                    else e.line("  const double den%d = M_(2.0, %s), inv%d = D_(1.0, den%d); const bool ok%d = tb_midrange(den%d);", opi, r, opi, opi, opi, opi);
                } else {
                    const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                    Terms t;
                    for (int i = 1; i <= kend; i++) term(t, C(i), C(k - i));
                    std::string s = sum(e, t);
                    if (fast) {
                        if (kodd) e.line("  const double %s = fma(-2.0, %s, %s) * inv%d;", r, s.c_str(), A(k).c_str(), opi);
                        else e.line("  const double %s = fma(-%s, %s, fma(-2.0, %s, %s)) * inv%d;", r, C(kend + 1).c_str(), C(kend + 1).c_str(), s.c_str(), A(k).c_str(), opi);
                    } else {
                        std::string aux = fmt("B_(%s, M_(2.0, %s))", A(k).c_str(), s.c_str());
                        if (!kodd) aux = fmt("B_(%s, M_(%s, %s))", aux.c_str(), C(kend + 1).c_str(), C(kend + 1).c_str());
                        e.line("  const double %s = tb_div_y(%s, den%d, inv%d, ok%d);", r, aux.c_str(), opi, opi, opi);
                    }
                }
                break;
            case TB_OP_POW:
                if (k0) {
                    e.line("  const double %s = tb_pow(%s, %s);", r, A(0).c_str(), dlit(pw).c_str());
                    if (fast) e.line("  const double inv%d = 1.0 / %s;", opi, A(0).c_str());
                } else {
                    Terms t;
                    for (int i = 0; i <= k - 1; i++) {
                        /* rr = pw*(k-i) - i, evaluated with the oracle's two IEEE operations */
                        volatile double m = pw * (double)(k - i);
                        volatile double rr = m - (double)i;
                        if (fast) {
                            std::string w = fmt("w%d", tmpc++);
                            e.line("  const double %s = %s * %s;", w.c_str(), dlit(rr).c_str(), C(i).c_str());
                            term(t, w, A(k - i));
                        } else term(t, fmt("M_(%s, %s)", dlit(rr).c_str(), C(i).c_str()), A(k - i));
                    }
                    std::string s = sum(e, t);
                    if (fast) e.line("  const double %s = %s * (inv%d * %s);", r, s.c_str(), opi, dlit(1.0 / k).c_str());
                    else e.line("  const double %s = D_(%s, M_(%s, %s));", r, s.c_str(), dlit((double)k).c_str(), A(0).c_str());
                }
                break;
            case TB_OP_EXP:
                if (k0) e.line("  const double %s = tb_exp(%s);", r, A(0).c_str());
                else {
                    Terms t;
                    for (int i = 0; i <= k - 1; i++) {
                        if (fast) {
                            std::string w = fmt("w%d", tmpc++);
                            e.line("  const double %s = %s * %s;", w.c_str(), dlit((double)(k - i)).c_str(), A(k - i).c_str());
                            term(t, w, C(i));
                        } else term(t, fmt("M_(%s, %s)", dlit((double)(k - i)).c_str(), A(k - i).c_str()), C(i));
                    }
                    std::string s = sum(e, t);
                    if (fast) e.line("  const double %s = %s * %s;", r, s.c_str(), dlit(1.0 / k).c_str());
                    else e.line("  const double %s = %s;", r, divc(s, k).c_str());
                }
                break;
            case TB_OP_LOG:
                if (k0) { e.line("  const double %s = tb_log(%s);", r, A(0).c_str()); if (fast) e.line("  const double inv%d = 1.0 / %s;", opi, A(0).c_str()); }
                else if (k == 1) {
                    if (fast) e.line("  const double %s = %s * inv%d;", r, A(1).c_str(), opi);
                    else e.line("  const double %s = D_(%s, %s);", r, A(1).c_str(), A(0).c_str());
                } else {
                    Terms t;
                    term(t, fmt("M_(%s, %s)", dlit((double)(k - 1)).c_str(), A(1).c_str()), C(k - 1));
                    for (int i = 2; i <= k - 1; i++) term(t, fmt("M_(%s, %s)", dlit((double)(k - i)).c_str(), A(i).c_str()), C(k - i));
                    std::string s = sum(e, t);
                    if (fast) e.line("  const double %s = fma(-%s, %s, %s) * inv%d;", r, s.c_str(), dlit(1.0 / k).c_str(), A(k).c_str(), opi);
                    else e.line("  const double %s = D_(B_(%s, %s), %s);", r, A(k).c_str(), divc(s, k).c_str(), A(0).c_str());
                }
                break;
            case TB_OP_SINCOS: {
                const std::string RC = vname(op.aux, k);
                auto Cc = [&](int idx) { return rd(op.aux, idx, k, defined); };
                if (k0) e.line("  double %s, %s; tb_sincos(%s, &%s, &%s);", r, RC.c_str(), A(0).c_str(), r, RC.c_str());
                else {
                    Terms ts, tc;
                    for (int i = 1; i <= k; i++) {
                        std::string x = A(i);
                        if (fast && x == "0.0") continue;
                        if (i >= 2) {
                            std::string w = fmt("w%d", tmpc++);
                            e.line("  const double %s = M_(%s, %s);", w.c_str(), dlit((double)i).c_str(), x.c_str());
                            x = w;
                        }
                        ts.push_back({x, Cc(k - i)});
                        tc.push_back({x, C(k - i)});
                    }
                    std::string s1 = sum(e, ts), s2 = sum(e, tc, true);
                    if (fast) {
                        e.line("  const double %s = %s * %s;", r, s1.c_str(), dlit(1.0 / k).c_str());
                        e.line("  const double %s = %s * %s;", RC.c_str(), s2.c_str(), dlit(1.0 / k).c_str());
                    } else {
                        e.line("  const double %s = %s;", r, divc(s1, k).c_str());
                        e.line("  const double %s = %s;", RC.c_str(), divc(s2, k).c_str());
                    }
                }
                store(e, op.aux, k, RC);
                defined.insert(op.aux);
                break;
            }
            default: err = "opcode not supported by the unrolled code generator"; break;
        }
        store(e, op.res, k, R);
        defined.insert(op.res);
    }

    /* parity-mode x / n for a small integer constant n: exact multiply for powers of two,
     * Markstein correction (tb_div_y) otherwise -- same bits as the IEEE division      */
    std::string divc(const std::string &x, int n) {
        if ((n & (n - 1)) == 0) return fmt("M_(%s, %s)", x.c_str(), dlit(1.0 / n).c_str());
        return fmt("tb_div_y(%s, %s, %s, true)", x.c_str(), dlit((double)n).c_str(), dlit(1.0 / n).c_str());
    }

    void store(Em &e, int32_t o, int k, const std::string &v) {
        auto it = ser.find(o);
        if (it == ser.end()) return;
        if (it->second.place == PL_REG) e.line("  J.r%d[%d] = %s;", it->second.id, k, v.c_str());
        else if (it->second.place == PL_SMEM) e.line("  S(%d, %d) = %s;", it->second.id, k, v.c_str());
    }

    /* jetcoeffs!: all orders, straight line */
    void emit_jets(Em &e) {
        const tb_ode_desc_t &d = ode.desc;
        e.line("__device__ __forceinline__ void tb_u_jets(TbJets &J, double *__restrict__ smt, const double *__restrict__ par,");
        e.line("                                          const double t0, const double *__restrict__ x0) {");
        for (int c = 0; c < d.dof; c++) {
            e.line("  const double vX%d_0 = x0[%d];", c, c);
            store(e, TB_OPERAND(TB_K_X, c), 0, fmt("vX%d_0", c));
        }
        for (int k = 0; k < P; k++) {
            e.line("  /* ---- order %d ---- */", k);
            if (k > 0 && !(co.flags & 8)) e.line("  asm volatile(\"\" ::: \"memory\");");
            std::set<int32_t> defined;
            for (int j = 0; j < d.nops; j++) emit_op(e, j, ode.ops[j], k, defined);
            for (int c = 0; c < d.dof; c++) { /* solcoeff!: x[k+1] = dx[k]/(k+1) */
                int32_t dx = TB_OPERAND(TB_K_DX, c);
                if (!defined.count(dx)) { err = "tape does not define dx[" + std::to_string(c) + "]"; return; }
                std::string v = fmt("vX%d_%d", c, k + 1);
                if (fast) e.line("  const double %s = %s * %s;", v.c_str(), vname(dx, k).c_str(), dlit(1.0 / (k + 1)).c_str());
                else e.line("  const double %s = %s;", v.c_str(), divc(vname(dx, k), k + 1).c_str());
                store(e, TB_OPERAND(TB_K_X, c), k + 1, v);
            }
        }
        e.line("}");
    }

    /* coefficient k of state jet c; resolves to a register, an LDS or a scaled dx
     * coefficient once the caller's loops over (c, k) are unrolled                */
    void emit_accessor(Em &e) {
        const tb_ode_desc_t &d = ode.desc;
        e.line("__device__ __forceinline__ double tb_u_x(const TbJets &J, const double *__restrict__ smt, const double *__restrict__ x0, const int c, const int k) {");
        e.line("  switch (c) {");
        for (int c = 0; c < d.dof; c++) {
            const SeriesInfo &si = ser[TB_OPERAND(TB_K_X, c)];
            if (si.place == PL_REG) e.line("    case %d: return J.r%d[k];", c, si.id);
            else if (si.place == PL_SMEM) e.line("    case %d: return S(%d, k);", c, si.id);
            else {
                const SeriesInfo &di = ser[TB_OPERAND(TB_K_DX, c)];
                std::string ref = di.place == PL_REG ? fmt("J.r%d[k - 1]", di.id) : fmt("S(%d, k - 1)", di.id);
                e.line("    case %d: return k == 0 ? x0[%d] : %s * (1.0 / (double)k);", c, c, ref.c_str());
            }
        }
        e.line("    default: return 0.0;");
        e.line("  }");
        e.line("}");
    }
};

} // namespace

int generate_unrolled(const Ode &ode, const tb_compile_opts_t &co, Layout &L, std::string &src, std::string &err) {
    const tb_ode_desc_t &d = ode.desc;
    if (co.kind != TB_KIND_PLAIN) { err = "the unrolled generator covers TB_KIND_PLAIN"; return TB_ERR_UNSUPPORTED; }
    if (co.order < 1 || co.order > 40) { err = "order must be in 1..40 for the unrolled generator"; return TB_ERR_UNSUPPORTED; }
    if (d.dof > 16 || d.nops > 200) { err = "tape too large to unroll"; return TB_ERR_UNSUPPORTED; }
    std::set<int32_t> seen;
    for (int j = 0; j < d.nops; j++) {
        const tb_op_t &op = ode.ops[j];
        if (!seen.insert(op.res).second || (op.op == TB_OP_SINCOS && !seen.insert(op.aux).second)) {
            err = "tape is not in single-assignment form";
            return TB_ERR_UNSUPPORTED;
        }
    }
    UGen g(ode, co, L);
    const int np1 = co.order + 1;
    int budget = co.reg_budget > 0 ? co.reg_budget : 3 * np1;
    if (co.reg_budget < 0) budget = 0;
    g.analyse(budget);
    L.variant = 2;
    L.nx = d.dof;
    L.nslot = g.nsm;
    L.slot.clear();
    const size_t per_traj = (size_t)std::max(1, g.nsm) * np1 * sizeof(double);
    const size_t smem_max = 227 * 1024;
    int bt = co.block_threads;
    if (bt <= 0) {
        bt = (int)(smem_max / per_traj) / 32 * 32;
        if (bt > 256) bt = 256;
    }
    if (bt < 32 || (bt % 32) || (size_t)bt * per_traj > smem_max) {
        err = "working set does not fit one warp in shared memory";
        return TB_ERR_UNSUPPORTED;
    }
    L.block_threads = bt;
    L.smem_bytes = (size_t)bt * per_traj;
    L.blocks_per_sm = 1;

    Em e;
    e.line("/* generated by libtaylorb200 tb_codegen_unrolled.cpp -- do not edit */");
    e.line("#define TB_VARIANT 2");
    e.line("#define TB_P %d", co.order);
    e.line("#define TB_DOF %d", d.dof);
    e.line("#define TB_NX %d", d.dof);
    e.line("#define TB_NPAR %d", d.npar);
    e.line("#define TB_NSLOT %d", g.nsm);
    e.line("#define TB_BT %d", bt);
    e.line("#define TB_MINB 1");
    e.line("#define TB_MODE %d", co.mode);
    e.line("#define TB_KIND %d", co.kind);
    e.line("#define TB_EV_NOPS 0");
    e.line("#define TB_JAC_NOPS 0");
    e.s += embedded_math_h();
    e.s += embedded_device_pre();
    e.line("/* placement: %d series in registers, %d in shared memory */", g.nreg, g.nsm);
    e.line("struct TbJets {");
    for (int i = 0; i < g.nreg; i++) e.line("  double r%d[%d];", i, np1);
    if (g.nreg == 0) e.line("  int unused;");
    e.line("};");
    g.emit_jets(e);
    g.emit_accessor(e);
    e.s += embedded_device_drv();
    if (!g.err.empty()) { err = g.err; return TB_ERR_ARG; }
    src.swap(e.s);
    return TB_OK;
}

} // namespace tb
