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

enum Place { PL_NONE = 0, PL_REG = 1, PL_SMEM = 2, PL_DERIVED = 3, PL_TMEM = 4 };

struct SeriesInfo {
    int place = PL_NONE;
    int id = -1;     /* register array index or shared-memory slot */
    double reads = 0; /* Cauchy-sum reads per step (placement weight) */
    int len = 0;      /* coefficients kept */
    int base = 0;     /* shared memory: coefficient k lives at double (base + k) of the thread's column */
    int trim = 0;     /* shared memory: coefficients 0..trim-1 are not stored (exactly recoverable, see analyse) */
    int32_t dx = 0;   /* trimmed state jet: its derivative series */
    int kreg = 0;     /* PL_REG: coefficients 0..kreg-1 live in registers, kreg..len-1 in shared memory (split series) */
    int tbase = 0;    /* PL_TMEM: coefficient k lives in columns 2*(tbase + k), 2*(tbase + k) + 1 of the thread's window */
};

struct UGen {
    const Ode &ode;
    const tb_compile_opts_t &co;
    Layout &L;
    const int P;
    const bool fast;
    std::map<int32_t, SeriesInfo> ser;
    std::string err;
    int nreg = 0, nsm = 0, ntm = 0;
    int tmem_doubles = 0; /* tensor-memory doubles per trajectory */
    int tmpc = 0;

    UGen(const Ode &o, const tb_compile_opts_t &c, Layout &l) : ode(o), co(c), L(l), P(c.order), fast(c.mode == TB_MODE_FAST) {}

    /* ---- analysis ---------------------------------------------------------- */
    void need(int32_t o, double w) {
        if (!u_is_series(o) || TB_OPERAND_KIND(o) == TB_K_TIME) return;
        ser[o].reads += w;
    }
    bool derive_x = true;
    bool split = false;   /* registers hold the LOW coefficients of every recurrence-read series (see analyse) */
    bool tmem_tails = false; /* TB_CF_TMEM_TAILS */
    int top_trim = 0;     /* register series keep coefficients 0..len-1-top_trim in registers, the rest in shared memory */
    bool trim_x = false;  /* parity: do not store x[0], x[1], x[2] of state jets that only Horner / stepsize read */
    int smem_doubles = 0; /* shared-memory doubles per trajectory */
    void analyse(int reg_budget, int tmem_budget) {
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
        /* state jets: stored unless no recurrence reads them and dx is stored: then x[k] = dx[k-1]/k is
         * recomputed where Horner / stepsize need it (same expression, same bits), which frees
         * shared memory for more resident trajectories                                           */
        for (int c = 0; c < d.dof; c++) {
            int32_t x = TB_OPERAND(TB_K_X, c), dx = TB_OPERAND(TB_K_DX, c);
            bool cauchy = ser.count(x) && ser[x].reads > 0;
            bool dx_stored = ser.count(dx) && ser[dx].reads > 0;
            if (!cauchy && dx_stored && derive_x) { ser[x].place = PL_DERIVED; ser[x].id = c; }
            else ser[x].reads += 0.0; /* creates the entry: stored */
        }
        /* placement by read count */
        std::vector<std::pair<double, int32_t>> order;
        for (auto &kv : ser)
            if (kv.second.place != PL_DERIVED) order.push_back({-kv.second.reads, kv.first});
        std::sort(order.begin(), order.end());
        /* Tensor memory as a third storage tier (B200: 256 KB per SM, idle in a kernel without tcgen05.mma): a series
         * that only the recurrences inside the jet body read (a temporary or a dx jet, not a state jet: Horner, the
         * step size and the output code never touch it) may live in the thread's private TMEM window, written with
         * tcgen05.st and read with tcgen05.ld (32x32b.x2: lane = thread, two 32-bit columns per double).  That frees
         * shared memory (more resident trajectories) or registers (room for ptxas to hoist loads over the 8-cycle
         * dependent FP64 chains).  The jet body must then run warp-converged (.sync.aligned), see tb_device_drv.cuh. */
        std::set<int32_t> no_tmem;
        for (int c = 0; c < d.dof; c++)
            if (ser[TB_OPERAND(TB_K_X, c)].place == PL_DERIVED) no_tmem.insert(TB_OPERAND(TB_K_DX, c));
        int used = 0, tused = 0;
        for (auto &pr : order) {
            SeriesInfo &si = ser[pr.second];
            si.len = P + 1;
            const bool state = TB_OPERAND_KIND(pr.second) == TB_K_X;
            if (si.reads > 0 && used + si.len <= reg_budget) {
                /* top_trim: the highest coefficients of a register series are read by a handful of terms per step
                 * (coefficient j only at orders >= j); keeping them in shared memory costs a few loads and hands the
                 * registers to ptxas, whose schedule of the dependent DMUL -> DADD chains is register-starved */
                si.place = PL_REG; si.id = nreg++; used += si.len; si.kreg = std::max(1, si.len - top_trim);
            } else if (!tmem_tails && si.reads > 0 && !state && !no_tmem.count(pr.second) && tused + si.len <= tmem_budget) {
                si.place = PL_TMEM; si.id = ntm++; tused += si.len;
            } else if (tmem_tails && state && si.reads <= 0 && tused + si.len <= tmem_budget) {
                /* TB_CF_TMEM_TAILS: the state jets that only Horner and the step size read (written once and read
                 * once per step) go to tensor memory; shared memory is left to the recurrence-read series and to the
                 * top coefficients of the register series, which is what lets more than 256 threads share an SM */
                si.place = PL_TMEM; si.id = ntm++; tused += si.len;
            } else { si.place = PL_SMEM; si.id = nsm++; }
        }
        if (split) {
            /* Split placement.  Coefficient j of a series read by recurrences is read P - j times per step, so the
             * top coefficients are cold: the register budget is spread over the LOW coefficients of ALL such series
             * and only their high coefficients go to shared memory.  Same registers, ~3x fewer LDS than keeping some
             * series whole in registers and the others whole in shared memory.                                      */
            int nread = 0;
            for (auto &pr : order) if (ser[pr.second].reads > 0) nread++;
            if (nread > 0) {
                int K = std::min(P + 1, std::max(1, reg_budget / nread));
                nreg = 0; nsm = 0;
                for (auto &pr : order) {
                    SeriesInfo &si = ser[pr.second];
                    if (si.reads > 0) { si.place = PL_REG; si.id = nreg++; si.kreg = std::min(K, si.len); }
                    else { si.place = PL_SMEM; si.id = nsm++; si.kreg = 0; }
                }
            }
        }
        /* trimmed state jets: x[0] is the state itself, x[1] = dx[0]/1 and x[2] = dx[1]/2 are exact scalings of
         * stored dx coefficients, so a state jet that no recurrence reads keeps only x[3..P] in shared memory.
         * For Kepler-28 that is what lets 256 instead of 224 trajectories share one SM.                       */
        if (trim_x && P >= 3)
            for (int c = 0; c < d.dof; c++) {
                int32_t x = TB_OPERAND(TB_K_X, c), dx = TB_OPERAND(TB_K_DX, c);
                SeriesInfo &si = ser[x];
                if ((si.place != PL_SMEM && si.place != PL_TMEM) || si.reads > 0 || !ser.count(dx) || ser[dx].place == PL_NONE || ser[dx].place == PL_DERIVED || ser[dx].place == PL_TMEM) continue;
                if (ser[dx].reads <= 0) continue;
                si.trim = 3;
                si.dx = dx;
            }
        /* shared-memory layout: untrimmed series first (slot * (P+1), as S(slot, k)), trimmed ones packed after */
        int run = 0;
        for (int pass = 0; pass < 2; pass++)
            for (auto &kv : ser) {
                SeriesInfo &si = kv.second;
                if (si.place != PL_SMEM || (si.trim != 0) != (pass == 1)) continue;
                si.base = run - si.trim;
                run += si.len - si.trim;
            }
        for (auto &kv : ser) { /* high coefficients of split register series */
            SeriesInfo &si = kv.second;
            if (si.place != PL_REG || si.kreg >= si.len) continue;
            si.base = run - si.kreg;
            run += si.len - si.kreg;
        }
        smem_doubles = run;
        int trun = 0;
        for (auto &kv : ser) {
            SeriesInfo &si = kv.second;
            if (si.place != PL_TMEM) continue;
            si.tbase = trun - si.trim;
            trun += si.len - si.trim;
        }
        tmem_doubles = trun;
    }
    std::string sref(const SeriesInfo &si, const std::string &k) const { return fmt("SB(%d, %s)", si.base, k.c_str()); }
    std::string sref(const SeriesInfo &si, int k) const { return fmt("SB(%d, %d)", si.base, k); }
    /* coefficient `k` (an expression that folds to a constant once the caller is unrolled) of a stored series */
    std::string dynref(const SeriesInfo &si, const std::string &k) const {
        if (si.place != PL_REG) return sref(si, k);
        if (si.kreg >= si.len) return fmt("J.r%d[%s]", si.id, k.c_str());
        return fmt("(%s < %d ? J.r%d[%s < %d ? %s : 0] : %s)", k.c_str(), si.kreg, si.id, k.c_str(), si.kreg, k.c_str(), sref(si, k).c_str());
    }

    /* ---- values ---------------------------------------------------------------
     * A coefficient (series o, index idx) is referred to by the NAME of the local
     * that computed it while idx >= cur (the lowest order with unfinished tape
     * entries: such values are still in registers), and re-read from its storage
     * (register array or one LDS.64) afterwards.  Loads are cached per emission
     * group, so a coefficient shared by the chains of a group is read once.      */
    struct Ref { int32_t o; int idx; };
    typedef std::pair<int32_t, int> Key;
    std::set<Key> computed;
    std::map<Key, std::string> loaded;
    std::map<std::pair<std::string, std::string>, std::string> scaled;
    int cur = 0;
    bool tm_dirty = false;

    static std::string vname(int32_t o, int k) {
        int kind = TB_OPERAND_KIND(o), i = TB_OPERAND_INDEX(o);
        return fmt("%s%d_%d", kind == TB_K_DX ? "vD" : (kind == TB_K_TMP ? "vT" : "vX"), i, k);
    }
    bool avail(const Ref &r) const {
        if (TB_OPERAND_KIND(r.o) == TB_K_TIME) return true;
        return computed.count(Key(r.o, r.idx)) != 0;
    }
    std::string val(Em &e, const Ref &r) {
        if (TB_OPERAND_KIND(r.o) == TB_K_TIME) return r.idx == 0 ? "t0" : (r.idx == 1 ? "1.0" : "0.0");
        if (!computed.count(Key(r.o, r.idx))) { err = "internal: coefficient read before it is computed"; return "0.0"; }
        if (r.idx >= cur) return vname(r.o, r.idx);
        auto it = ser.find(r.o);
        if (it == ser.end() || it->second.place == PL_NONE || it->second.place == PL_DERIVED) {
            err = "internal: series operand not stored";
            return "0.0";
        }
        if (it->second.place == PL_REG && r.idx < it->second.kreg) return fmt("J.r%d[%d]", it->second.id, r.idx);
        auto lt = loaded.find(Key(r.o, r.idx));
        if (lt != loaded.end()) return lt->second;
        std::string n = fmt("l%d", tmpc++);
        if (r.idx < it->second.trim) { err = "internal: trimmed coefficient read by a recurrence"; return "0.0"; }
        if (it->second.place == PL_TMEM) {
            /* tcgen05.wait::st before the first tensor-memory load that follows a tensor-memory store */
            if (tm_dirty) { e.line("  TB_TM_FENCE();"); tm_dirty = false; }
            e.line("  const double %s = TB_TM_LD(%d);", n.c_str(), 2 * (it->second.tbase + r.idx));
        }
        else e.line("  const double %s = %s;", n.c_str(), sref(it->second, r.idx).c_str());
        loaded[Key(r.o, r.idx)] = n;
        return n;
    }
    std::string num(int32_t o, const tb_op_t &op) const {
        if (TB_OPERAND_KIND(o) == TB_K_PAR) return fmt("par[%d]", TB_OPERAND_INDEX(o));
        return dlit(op.imm);
    }
    void define(Em &e, int32_t o, int k, const std::string &name) {
        store(e, o, k, name);
        computed.insert(Key(o, k));
    }
    static int eff_op(const tb_op_t &op) {
        if (op.op != TB_OP_POW) return op.op;
        if (op.imm == 0.0) return TB_OP_ONE;
        if (op.imm == 1.0) return TB_OP_IDENTITY;
        if (op.imm == 2.0) return TB_OP_SQR;
        if (op.imm == 0.5) return TB_OP_SQRT;
        return TB_OP_POW;
    }
    /* does entry `op` at order k run a recurrence sum? */
    static bool is_sum_node(const tb_op_t &op, int k) {
        if (k == 0) return false;
        const bool sa = u_is_num(op.a), sb = u_is_num(op.b);
        switch (eff_op(op)) {
            case TB_OP_MUL: return !sa && !sb;
            case TB_OP_DIV: return !sb;
            case TB_OP_SQR: case TB_OP_SQRT: case TB_OP_POW: case TB_OP_EXP: case TB_OP_SINCOS: return true;
            case TB_OP_LOG: return k >= 2;
            default: return false;
        }
    }
    /* coefficients entry `op` reads at order k (conservative full ranges) */
    void deps(const tb_op_t &op, int k, std::vector<Ref> &out) const {
        const bool sa = u_is_num(op.a), sb = u_is_num(op.b);
        auto range = [&](int32_t o, int lo, int hi) { if (u_is_series(o)) for (int i = lo; i <= hi; i++) out.push_back({o, i}); };
        switch (eff_op(op)) {
            case TB_OP_IDENTITY: case TB_OP_NEG: if (!sa) range(op.a, k, k); break;
            case TB_OP_ZERO: case TB_OP_ONE: break;
            case TB_OP_ADD: case TB_OP_SUB: if (!sa) range(op.a, k, k); if (!sb) range(op.b, k, k); break;
            case TB_OP_MUL:
                if (sa && sb) break;
                if (sa) range(op.b, k, k); else if (sb) range(op.a, k, k);
                else { range(op.a, 0, k); range(op.b, 0, k); }
                break;
            case TB_OP_SQR: range(op.a, 0, k); break;
            case TB_OP_DIV:
                if (sa && sb) break;
                if (sb) { range(op.a, k, k); break; }
                range(op.b, 0, k); if (!sa) range(op.a, k, k); range(op.res, 0, k - 1);
                break;
            case TB_OP_SQRT: range(op.a, k, k); range(op.res, 0, k - 1); break;
            case TB_OP_POW: case TB_OP_EXP: case TB_OP_LOG: range(op.a, 0, k); range(op.res, 0, k - 1); break;
            case TB_OP_SINCOS: range(op.a, 0, k); range(op.res, 0, k - 1); range(op.aux, 0, k - 1); break;
            default: break;
        }
    }
    /* ready: every dependency computed, ignoring coefficients `skip_k` of the entry's own outputs */
    bool node_ready(const tb_op_t &op, int k, int skip_k = -1) const {
        std::vector<Ref> d;
        deps(op, k, d);
        for (const Ref &r : d) {
            if (skip_k >= 0 && r.idx == skip_k && (r.o == op.res || (op.op == TB_OP_SINCOS && r.o == op.aux))) continue;
            if (!avail(r)) return false;
        }
        return true;
    }

    /* ---- entries without a recurrence sum: order 0 of everything, linear ops, log at k = 1 ---- */
    void emit_cheap(Em &e, int opi, const tb_op_t &op, int k) {
        const bool sa = u_is_num(op.a), sb = u_is_num(op.b);
        const bool k0 = k == 0;
        const std::string R = vname(op.res, k);
        const char *r = R.c_str();
        const double pw = op.imm;
        /* operand strings are resolved first: val() may emit a load line */
        auto A = [&](int idx) { return sa ? num(op.a, op) : val(e, {op.a, idx}); };
        auto B = [&](int idx) { return sb ? num(op.b, op) : val(e, {op.b, idx}); };
        switch (eff_op(op)) {
            case TB_OP_IDENTITY:
                if (sa) e.line("  const double %s = %s;", r, k0 ? A(0).c_str() : "0.0");
                else { std::string a = A(k); e.line("  const double %s = %s;", r, a.c_str()); }
                break;
            case TB_OP_NEG: { std::string a = A(k); e.line("  const double %s = N_(%s);", r, a.c_str()); break; }
            case TB_OP_ZERO: e.line("  const double %s = 0.0;", r); break;
            case TB_OP_ONE: e.line("  const double %s = %s;", r, k0 ? "1.0" : "0.0"); break;
            case TB_OP_ADD:
            case TB_OP_SUB: {
                const bool add = eff_op(op) == TB_OP_ADD;
                const char *F = add ? "A_" : "B_";
                if (k0 || (!sa && !sb)) {
                    if (sa && sb && !k0) { e.line("  const double %s = 0.0;", r); break; }
                    std::string a = A(k), b = B(k);
                    e.line("  const double %s = %s(%s, %s);", r, F, a.c_str(), b.c_str());
                } else if (sa && sb) e.line("  const double %s = 0.0;", r);
                else if (sa) {
                    std::string b = B(k);
                    if (add) e.line("  const double %s = %s;", r, b.c_str());
                    else e.line("  const double %s = N_(%s);", r, b.c_str());
                } else { std::string a = A(k); e.line("  const double %s = %s;", r, a.c_str()); }
                break;
            }
            case TB_OP_MUL:
                if (sa && sb && !k0) e.line("  const double %s = 0.0;", r);
                else { std::string a = A(k), b = B(k); e.line("  const double %s = M_(%s, %s);", r, a.c_str(), b.c_str()); }
                break;
            case TB_OP_SQR: { std::string a = A(0); e.line("  const double %s = M_(%s, %s);", r, a.c_str(), a.c_str()); break; }
            case TB_OP_DIV:
                if (sa && sb) {
                    if (k0) { std::string a = A(0), b = B(0); e.line("  const double %s = D_(%s, %s);", r, a.c_str(), b.c_str()); }
                    else e.line("  const double %s = 0.0;", r);
                } else if (sb) {
                    std::string a = A(k), b = B(k);
                    if (fast) { if (k0) e.line("  const double inv%d = 1.0 / %s;", opi, b.c_str()); e.line("  const double %s = %s * inv%d;", r, a.c_str(), opi); }
                    else e.line("  const double %s = %s(%s, %s);", r, k0 ? "D_" : "DIVQ_", a.c_str(), b.c_str());
                } else { /* k == 0 of a series divisor */
                    std::string a = A(0), b = B(0);
                    if (fast) { e.line("  const double inv%d = 1.0 / %s;", opi, b.c_str()); e.line("  const double %s = %s * inv%d;", r, a.c_str(), opi); }
                    else {
                        e.line("  const double den%d = %s, inv%d = D_(1.0, den%d); const bool ok%d = tb_midrange(den%d); DENCHK_(den%d);", opi, b.c_str(), opi, opi, opi, opi, opi);
                        e.line("  const double %s = D_(%s, den%d);", r, a.c_str(), opi);
                    }
                }
                break;
            case TB_OP_SQRT: {
                std::string a = A(0);
                e.line("  const double %s = Q_(%s);", r, a.c_str());
                if (fast) e.line("  const double inv%d = 0.5 / %s;", opi, r);
                else e.line("  const double den%d = M_(2.0, %s), inv%d = D_(1.0, den%d); const bool ok%d = tb_midrange(den%d); DENCHK_(den%d);", opi, r, opi, opi, opi, opi, opi);
                break;
            }
            case TB_OP_POW: {
                std::string a = A(0);
                e.line("  const double %s = tb_pow(%s, %s);", r, a.c_str(), dlit(pw).c_str());
                if (fast) e.line("  const double inv%d = 1.0 / %s;", opi, a.c_str());
                break;
            }
            case TB_OP_EXP: { std::string a = A(0); e.line("  const double %s = tb_exp(%s);", r, a.c_str()); break; }
            case TB_OP_LOG:
                if (k0) {
                    std::string a = A(0);
                    e.line("  const double %s = tb_log(%s);", r, a.c_str());
                    if (fast) e.line("  const double inv%d = 1.0 / %s;", opi, a.c_str());
                } else { /* k == 1 */
                    std::string a1 = A(1), a0 = A(0);
                    if (fast) e.line("  const double %s = %s * inv%d;", r, a1.c_str(), opi);
                    else e.line("  const double %s = D_(%s, %s);", r, a1.c_str(), a0.c_str());
                }
                break;
            case TB_OP_SINCOS: {
                const std::string RC = vname(op.aux, k);
                std::string a = A(0);
                e.line("  double %s, %s; tb_sincos(%s, &%s, &%s);", r, RC.c_str(), a.c_str(), r, RC.c_str());
                define(e, op.aux, k, RC);
                break;
            }
            default: err = "opcode not supported by the unrolled code generator"; break;
        }
        define(e, op.res, k, R);
    }

    /* ---- recurrence sums --------------------------------------------------------
     * A chain is one left-to-right sum of products (coef*u)*v.  Parity mode emits
     * acc = acc + (coef*u)*v with unfused operations in the oracle's order; fast
     * mode uses FMA and `nacc` partial accumulators.  The chains of a group are
     * emitted round-robin, one term each, so that (a) the two orders of a pair
     * share every loaded coefficient and (b) independent add chains interleave.  */
    struct Term { bool has_coef; double coef; Ref u, v; };
    struct Chain {
        std::vector<Term> terms;
        bool neg = false;
        size_t pos = 0;
        std::string acc;
        int nacc = 1;
    };
    struct SNode {
        int j = 0, k = 0;
        std::vector<Chain> ch;
        bool fin = false;
    };

    void push_term(Chain &c, const Ref &u, const Ref &v, bool has_coef = false, double coef = 0.0) {
        auto tz = [&](const Ref &r) { return TB_OPERAND_KIND(r.o) == TB_K_TIME && r.idx >= 2; };
        if (fast && (tz(u) || tz(v))) return; /* structural zeros of the time series */
        c.terms.push_back({has_coef, coef, u, v});
    }
    void build_chains(SNode &n) {
        const tb_op_t &op = ode.ops[n.j];
        const int k = n.k;
        n.ch.assign(1, Chain());
        Chain &c = n.ch[0];
        switch (eff_op(op)) {
            case TB_OP_MUL: for (int i = 0; i <= k; i++) push_term(c, {op.a, i}, {op.b, k - i}); break;
            case TB_OP_SQR: { const int kodd = k & 1, kend = (k - 2 + kodd) / 2; for (int i = 0; i <= kend; i++) push_term(c, {op.a, i}, {op.a, k - i}); break; }
            case TB_OP_DIV: for (int i = 0; i <= k - 1; i++) push_term(c, {op.res, i}, {op.b, k - i}); break;
            case TB_OP_SQRT: { const int kodd = k & 1, kend = (k - 2 + kodd) / 2; for (int i = 1; i <= kend; i++) push_term(c, {op.res, i}, {op.res, k - i}); break; }
            case TB_OP_POW:
                for (int i = 0; i <= k - 1; i++) { /* rr = pw*(k-i) - i with the oracle's two IEEE operations */
                    volatile double m = op.imm * (double)(k - i);
                    volatile double rr = m - (double)i;
                    push_term(c, {op.res, i}, {op.a, k - i}, true, rr);
                }
                break;
            case TB_OP_EXP: for (int i = 0; i <= k - 1; i++) push_term(c, {op.a, k - i}, {op.res, i}, true, (double)(k - i)); break;
            case TB_OP_LOG:
                push_term(c, {op.a, 1}, {op.res, k - 1}, true, (double)(k - 1));
                for (int i = 2; i <= k - 1; i++) push_term(c, {op.a, i}, {op.res, k - i}, true, (double)(k - i));
                break;
            case TB_OP_SINCOS: {
                n.ch.push_back(Chain());
                Chain &s = n.ch[0], &cc = n.ch[1];
                cc.neg = true;
                for (int i = 1; i <= k; i++) {
                    push_term(s, {op.a, i}, {op.aux, k - i}, i >= 2, (double)i);
                    push_term(cc, {op.a, i}, {op.res, k - i}, i >= 2, (double)i);
                }
                break;
            }
            default: break;
        }
    }
    /* coef*u, computed once per group (the two chains of sincos! share it) */
    std::string scaled_val(Em &e, const Term &t) {
        std::string u = val(e, t.u);
        if (!t.has_coef) return u;
        std::string c = dlit(t.coef);
        auto key = std::make_pair(c, u);
        auto it = scaled.find(key);
        if (it != scaled.end()) return it->second;
        std::string w = fmt("w%d", tmpc++);
        e.line("  const double %s = M_(%s, %s);", w.c_str(), c.c_str(), u.c_str());
        scaled[key] = w;
        return w;
    }
    void emit_term(Em &e, Chain &c, const Term &t) {
        std::string x = scaled_val(e, t), y = val(e, t.v);
        const size_t i = c.pos;
        if (c.acc.empty()) c.acc = fmt("a%d", tmpc++);
        const char *a = c.acc.c_str();
        if (!fast) {
            if (i == 0) e.line("  double %s = M_(%s%s%s, %s);", a, c.neg ? "N_(" : "", x.c_str(), c.neg ? ")" : "", y.c_str());
            else if (mulahead > 0) {
                /* product and add on separate lines: hoist_lines() moves the product `mulahead` lines up so that the
                 * DADD finds it finished (ptxas keeps the source order and would otherwise issue DMUL t; DADD acc, t
                 * back to back through one temporary: 8 cycles of dependent-issue latency per term)             */
                std::string pn = fmt("p%d", tmpc++);
                e.line("  const double %s = M_(%s, %s);", pn.c_str(), x.c_str(), y.c_str());
                e.line("  %s = %s(%s, %s);", a, c.neg ? "B_" : "A_", a, pn.c_str());
            } else e.line("  %s = %s(%s, M_(%s, %s));", a, c.neg ? "B_" : "A_", a, x.c_str(), y.c_str());
        } else {
            const int j = (int)(i % (size_t)c.nacc);
            if (i < (size_t)c.nacc) e.line("  double %s_%d = %s * %s;", a, j, x.c_str(), y.c_str());
            else e.line("  %s_%d = fma(%s, %s, %s_%d);", a, j, x.c_str(), y.c_str(), a, j);
        }
    }
    /* name (or literal) holding the finished sum */
    std::string chain_sum(Em &e, Chain &c) {
        if (c.terms.empty()) return "0.0";
        if (!fast) return c.acc;
        const int n = (int)std::min<size_t>((size_t)c.nacc, c.terms.size());
        const char *a = c.acc.c_str();
        std::string comb;
        if (n == 1) comb = fmt("%s_0", a);
        else if (n == 2) comb = fmt("%s_0 + %s_1", a, a);
        else if (n == 3) comb = fmt("(%s_0 + %s_1) + %s_2", a, a, a);
        else comb = fmt("(%s_0 + %s_1) + (%s_2 + %s_3)", a, a, a, a);
        std::string s = fmt("s%d", tmpc++);
        e.line("  const double %s = %s(%s);", s.c_str(), c.neg ? "-" : "", comb.c_str());
        return s;
    }
    bool fin_ready(const SNode &n) const {
        const tb_op_t &op = ode.ops[n.j];
        const int k = n.k, o = eff_op(op);
        if ((o == TB_OP_DIV && !u_is_num(op.a)) || o == TB_OP_SQRT || o == TB_OP_LOG) if (!avail({op.a, k})) return false;
        if (o == TB_OP_SQR && !(k & 1) && !avail({op.a, k / 2})) return false;
        if (o == TB_OP_SQRT && !(k & 1) && !avail({op.res, (k - 2) / 2 + 1})) return false;
        return true;
    }
    void finalize(Em &e, SNode &n) {
        const tb_op_t &op = ode.ops[n.j];
        const int k = n.k, opi = n.j, o = eff_op(op);
        const bool sa = u_is_num(op.a);
        const std::string R = vname(op.res, k);
        const char *r = R.c_str();
        std::string s = chain_sum(e, n.ch[0]);
        auto A = [&](int idx) { return val(e, {op.a, idx}); };
        switch (o) {
            case TB_OP_MUL: e.line("  const double %s = %s;", r, s.c_str()); break;
            case TB_OP_SQR:
                if (k & 1) e.line("  const double %s = M_(2.0, %s);", r, s.c_str());
                else {
                    std::string h = A(k / 2);
                    if (fast) e.line("  const double %s = fma(%s, %s, 2.0 * %s);", r, h.c_str(), h.c_str(), s.c_str());
                    else e.line("  const double %s = A_(M_(2.0, %s), M_(%s, %s));", r, s.c_str(), h.c_str(), h.c_str());
                }
                break;
            case TB_OP_DIV:
                if (fast) {
                    if (sa) e.line("  const double %s = -%s * inv%d;", r, s.c_str(), opi);
                    else { std::string a = A(k); e.line("  const double %s = (%s - %s) * inv%d;", r, a.c_str(), s.c_str(), opi); }
                } else {
                    if (sa) e.line("  const double %s = DIVY_(N_(%s), den%d, inv%d, ok%d);", r, s.c_str(), opi, opi, opi);
                    else { std::string a = A(k); e.line("  const double %s = DIVY_(B_(%s, %s), den%d, inv%d, ok%d);", r, a.c_str(), s.c_str(), opi, opi, opi); }
                }
                break;
            case TB_OP_SQRT: {
                std::string a = A(k);
                const bool even = !(k & 1);
                std::string c = even ? val(e, {op.res, (k - 2) / 2 + 1}) : std::string();
                if (fast) {
                    if (!even) e.line("  const double %s = fma(-2.0, %s, %s) * inv%d;", r, s.c_str(), a.c_str(), opi);
                    else e.line("  const double %s = fma(-%s, %s, fma(-2.0, %s, %s)) * inv%d;", r, c.c_str(), c.c_str(), s.c_str(), a.c_str(), opi);
                } else {
                    std::string aux = fmt("B_(%s, M_(2.0, %s))", a.c_str(), s.c_str());
                    if (even) aux = fmt("B_(%s, M_(%s, %s))", aux.c_str(), c.c_str(), c.c_str());
                    e.line("  const double %s = DIVY_(%s, den%d, inv%d, ok%d);", r, aux.c_str(), opi, opi, opi);
                }
                break;
            }
            case TB_OP_POW:
                if (fast) e.line("  const double %s = %s * (inv%d * %s);", r, s.c_str(), opi, dlit(1.0 / k).c_str());
                else if (pow_recip) /* s / (k a[0]) from the reciprocal prepared at the start of the group (emit_group): 3 dependent operations */
                    e.line("  const double %s = DIVY_(%s, denP%d_%d, invP%d_%d, okP%d_%d);", r, s.c_str(), opi, k, opi, k, opi, k);
                else { std::string a0 = A(0); e.line("  const double %s = DIVQ_(%s, M_(%s, %s));", r, s.c_str(), dlit((double)k).c_str(), a0.c_str()); }
                break;
            case TB_OP_EXP:
                if (fast) e.line("  const double %s = %s * %s;", r, s.c_str(), dlit(1.0 / k).c_str());
                else e.line("  const double %s = %s;", r, divc(s, k).c_str());
                break;
            case TB_OP_LOG: {
                std::string a = A(k);
                if (fast) e.line("  const double %s = fma(-%s, %s, %s) * inv%d;", r, s.c_str(), dlit(1.0 / k).c_str(), a.c_str(), opi);
                else { std::string a0 = A(0); e.line("  const double %s = DIVQ_(B_(%s, %s), %s);", r, a.c_str(), divc(s, k).c_str(), a0.c_str()); }
                break;
            }
            case TB_OP_SINCOS: {
                const std::string RC = vname(op.aux, k);
                std::string s2 = chain_sum(e, n.ch[1]);
                if (fast) {
                    e.line("  const double %s = %s * %s;", r, s.c_str(), dlit(1.0 / k).c_str());
                    e.line("  const double %s = %s * %s;", RC.c_str(), s2.c_str(), dlit(1.0 / k).c_str());
                } else {
                    e.line("  const double %s = %s;", r, divc(s, k).c_str());
                    e.line("  const double %s = %s;", RC.c_str(), divc(s2, k).c_str());
                }
                define(e, op.aux, k, RC);
                break;
            }
            default: err = "internal: not a sum entry"; break;
        }
        define(e, op.res, k, R);
        n.fin = true;
    }

    /* Software prefetch.  A storage load is emitted where its first use is, and ptxas keeps a volatile load where the
     * source has it: the consumer then eats the whole LDS latency (~30 cycles against ~8 cycles of FP64 issue per
     * term with two warps per scheduler; ncu: the DMUL after each LDS holds 2/3 of the samples of such a chain).
     * Every load of a group reads a coefficient stored by an EARLIER group (val() loads only indices below `cur`), so
     * it may move up to the start of the group: each load line moves `prefetch` lines up, order among loads kept.  */
    bool pow_recip = false; /* parity: pow! quotients from a reciprocal prepared ahead of the sum (TB_CF_POW_RECIP, opt-in: measured slower) */
    int prefetch = 0;  /* lines a storage load moves up (TB_CF_PREFETCH) */
    int mulahead = 0;  /* lines a product moves up (TB_CF_MULAHEAD) */
    static bool ident_char(char ch) { return (ch >= 'a' && ch <= 'z') || (ch >= 'A' && ch <= 'Z') || (ch >= '0' && ch <= '9') || ch == '_'; }
    void hoist_lines(Em &e, size_t from) {
        if ((prefetch <= 0 && mulahead <= 0) || from >= e.s.size()) return;
        std::vector<std::string> lines;
        size_t p = from;
        while (p < e.s.size()) {
            size_t q = e.s.find('\n', p);
            if (q == std::string::npos) q = e.s.size();
            lines.push_back(e.s.substr(p, q - p));
            p = q + 1;
        }
        auto starts = [](const std::string &l, const char *pre) { return l.compare(0, strlen(pre), pre) == 0; };
        auto is_load = [&](const std::string &l) {
            return starts(l, "  const double l") && (l.find("= SB(") != std::string::npos || l.find("= TB_TM_LD(") != std::string::npos);
        };
        auto is_prod = [&](const std::string &l) { return starts(l, "  const double p") || starts(l, "  const double w"); };
        /* name defined by a line ("  [const] double NAME = ...", "  double A, B; ..." is left alone) */
        auto defined = [&](const std::string &l) -> std::string {
            size_t d = l.find("double ");
            if (d == std::string::npos || d > 8) return std::string();
            size_t b = d + 7, q = b;
            while (q < l.size() && ident_char(l[q])) q++;
            return l.substr(b, q - b);
        };
        std::vector<std::string> out;
        std::map<std::string, size_t> defpos; /* name -> index in `out` (kept current across insertions) */
        size_t load_floor = 0;
        auto insert_at = [&](size_t pos, const std::string &l) {
            for (auto &kv : defpos) if (kv.second >= pos) kv.second++;
            out.insert(out.begin() + (long)pos, l);
            std::string nm = defined(l);
            if (!nm.empty()) defpos[nm] = pos;
        };
        for (const std::string &l : lines) {
            const bool ld = prefetch > 0 && is_load(l), pr = mulahead > 0 && is_prod(l);
            if (l.find("TB_TM_FENCE") != std::string::npos) { insert_at(out.size(), l); load_floor = out.size(); continue; }
            if (!ld && !pr) { insert_at(out.size(), l); continue; }
            const size_t dist = (size_t)(ld ? prefetch : mulahead);
            size_t pos = out.size() > dist ? out.size() - dist : 0;
            if (ld) pos = std::max(pos, load_floor);
            else { /* after the definition of everything the product reads */
                const size_t eq = l.find('=');
                size_t i = eq == std::string::npos ? 0 : eq + 1;
                while (i < l.size()) {
                    if (!ident_char(l[i]) || (l[i] >= '0' && l[i] <= '9')) { i++; continue; }
                    size_t j = i;
                    while (j < l.size() && ident_char(l[j])) j++;
                    auto it = defpos.find(l.substr(i, j - i));
                    if (it != defpos.end()) pos = std::max(pos, it->second + 1);
                    i = j;
                }
            }
            insert_at(pos, l);
            if (ld) load_floor = pos + 1;
        }
        e.s.resize(from);
        for (const std::string &l : out) { e.s += l; e.s += "\n"; }
    }

    /* emit a group of sum nodes, chains interleaved one term per sweep */
    void emit_group(Em &e, std::vector<SNode> &g) {
        const size_t group_start = e.s.size();
        loaded.clear();
        scaled.clear();
        size_t nchains = 0;
        for (SNode &n : g) { build_chains(n); nchains += n.ch.size(); }
        if (!fast && pow_recip)
            for (SNode &n : g) {
                /* pow!: c[k] = s / (k a[0]).  The divisor is known before the sum starts, so its correctly rounded
                 * reciprocal is computed here, off the critical path (it overlaps the add chain), and the quotient
                 * is finished with the Markstein correction: the full division (10 dependent operations between the
                 * end of this sum and the start of the division sums that wait for r[k]) leaves the chain.       */
                const tb_op_t &op = ode.ops[n.j];
                if (eff_op(op) != TB_OP_POW) continue;
                std::string a0 = val(e, {op.a, 0});
                e.line("  const double denP%d_%d = M_(%s, %s), invP%d_%d = D_(1.0, denP%d_%d); const bool okP%d_%d = tb_midrange(denP%d_%d); DENCHK_(denP%d_%d);",
                       n.j, n.k, dlit((double)n.k).c_str(), a0.c_str(), n.j, n.k, n.j, n.k, n.j, n.k, n.j, n.k, n.j, n.k);
            }
        for (SNode &n : g)
            for (Chain &c : n.ch) c.nacc = fast ? (nchains >= 4 ? 2 : 4) : 1;
        bool progress = true;
        while (progress) {
            progress = false;
            for (SNode &n : g) {
                if (n.fin) continue;
                bool all = true;
                for (Chain &c : n.ch) {
                    if (c.pos < c.terms.size()) {
                        const Term &t = c.terms[c.pos];
                        if (avail(t.u) && avail(t.v)) { emit_term(e, c, t); c.pos++; progress = true; }
                    }
                    if (c.pos < c.terms.size()) all = false;
                }
                if (all && fin_ready(n)) { finalize(e, n); progress = true; }
            }
        }
        for (SNode &n : g) if (!n.fin && err.empty()) err = "internal: recurrence group did not complete";
        hoist_lines(e, group_start);
    }
    /* parity-mode x / n for a small integer constant n: exact multiply for powers of two,
     * Markstein correction (tb_div_y) otherwise -- same bits as the IEEE division      */
    std::string divc(const std::string &x, int n) {
        if ((n & (n - 1)) == 0) return fmt("M_(%s, %s)", x.c_str(), dlit(1.0 / n).c_str());
        return fmt("DIVY_(%s, %s, %s, true)", x.c_str(), dlit((double)n).c_str(), dlit(1.0 / n).c_str());
    }

    void store(Em &e, int32_t o, int k, const std::string &v) {
        auto it = ser.find(o);
        if (it == ser.end()) return;
        if (it->second.place == PL_REG && k < it->second.kreg) e.line("  J.r%d[%d] = %s;", it->second.id, k, v.c_str());
        else if (it->second.place == PL_REG) e.line("  %s = %s;", sref(it->second, k).c_str(), v.c_str());
        else if (it->second.place == PL_TMEM) { if (k >= it->second.trim) { e.line("  TB_TM_ST(%d, %s);", 2 * (it->second.tbase + k), v.c_str()); tm_dirty = true; } }
        else if (it->second.place == PL_SMEM && k >= it->second.trim) e.line("  %s = %s;", sref(it->second, k).c_str(), v.c_str());
    }


    /* jetcoeffs!: dataflow-scheduled straight-line code.  Entries whose inputs are
     * known are emitted as early as possible; ready recurrence entries of the lowest
     * order form a group, and each takes its next order along as a PAIR when that
     * one only waits for the first's own result (x'' = f systems such as Kepler and
     * the pendulum: x[k+2] follows from order k).  Every value is computed by the
     * same expression as in order-by-order emission, so parity mode stays bitwise. */
    void emit_jets(Em &e) {
        const tb_ode_desc_t &d = ode.desc;
        const bool pairing = !(co.flags & 16), grouping = !(co.flags & 32);
        Em &out = e;
        Em body;
        {
        Em &e = body;
        tm_dirty = false; /* a step only loads coefficients it has stored itself: its first tensor-memory load follows a store */
        for (int c = 0; c < d.dof; c++) {
            e.line("  const double vX%d_0 = x0[%d];", c, c);
            define(e, TB_OPERAND(TB_K_X, c), 0, fmt("vX%d_0", c));
        }
        std::vector<std::vector<char>> done(d.nops, std::vector<char>(P, 0));
        std::vector<std::vector<char>> xdone(d.dof, std::vector<char>(P + 1, 0));
        long remaining = (long)d.nops * P + (long)d.dof * P;
        auto update_cur = [&]() {
            int c = P;
            for (int j = 0; j < d.nops; j++)
                for (int k = 0; k < P; k++) if (!done[j][k]) { c = std::min(c, k); break; }
            cur = c;
        };
        int last_sync_k = -1;
        while (remaining > 0 && err.empty()) {
            bool progress = true, any = false;
            while (progress) { /* entries without sums, and solcoeff!, to a fixed point */
                progress = false;
                for (int k = 0; k < P; k++) {
                    for (int j = 0; j < d.nops; j++) {
                        const tb_op_t &op = ode.ops[j];
                        if (done[j][k] || is_sum_node(op, k) || !node_ready(op, k)) continue;
                        emit_cheap(e, j, op, k);
                        done[j][k] = 1; remaining--; progress = any = true;
                    }
                    for (int c = 0; c < d.dof; c++) { /* solcoeff!: x[k+1] = dx[k]/(k+1) */
                        const int32_t dx = TB_OPERAND(TB_K_DX, c);
                        if (xdone[c][k + 1] || !computed.count(Key(dx, k))) continue;
                        std::string v = fmt("vX%d_%d", c, k + 1);
                        if (fast) e.line("  const double %s = %s * %s;", v.c_str(), vname(dx, k).c_str(), dlit(1.0 / (k + 1)).c_str());
                        else e.line("  const double %s = %s;", v.c_str(), divc(vname(dx, k), k + 1).c_str());
                        define(e, TB_OPERAND(TB_K_X, c), k + 1, v);
                        xdone[c][k + 1] = 1; remaining--; progress = any = true;
                    }
                }
            }
            update_cur();
            std::vector<SNode> g;
            int kmin = -1;
            for (int k = 1; k < P && kmin < 0; k++)
                for (int j = 0; j < d.nops; j++)
                    if (!done[j][k] && is_sum_node(ode.ops[j], k) && node_ready(ode.ops[j], k)) { kmin = k; break; }
            if (kmin >= 0) {
                for (int j = 0; j < d.nops; j++) {
                    const tb_op_t &op = ode.ops[j];
                    if (done[j][kmin] || !is_sum_node(op, kmin) || !node_ready(op, kmin)) continue;
                    SNode n; n.j = j; n.k = kmin;
                    g.push_back(n);
                    done[j][kmin] = 1; remaining--;
                    if (pairing && kmin + 1 < P && !done[j][kmin + 1] && is_sum_node(op, kmin + 1) && node_ready(op, kmin + 1, kmin)) {
                        SNode m; m.j = j; m.k = kmin + 1;
                        g.push_back(m);
                        done[j][kmin + 1] = 1; remaining--;
                    }
                    if (!grouping) break;
                }
                if ((co.flags & 128) && kmin != last_sync_k && (!(co.flags & 2048) || kmin % 4 == 1)) { e.line("  TB_ORDER_SYNC();"); last_sync_k = kmin; }
                e.line("  /* ---- order %d: %d recurrence chain group ---- */", kmin, (int)g.size());
                emit_group(e, g);
                e.line("  asm volatile(\"\" ::: \"memory\");");
                any = true;
            }
            if (!any) { err = "tape has a dependency cycle or an undefined operand (dx not defined?)"; break; }
        }
        }
        const char *sig = "(TbJets &J, double *__restrict__ smt, const unsigned tm, const double *__restrict__ par, const double t0, const double *__restrict__ x0)";
        const bool quick = !fast && !(co.flags & 64);
        if (quick) {
            /* parity mode, two copies of the body: the QUICK one divides with unguarded Markstein /
             * Newton sequences and folds every operand's exponent into one running maximum; when
             * that leaves the proven range (zero, subnormal-ish, huge, Inf, NaN operands) the step
             * is recomputed by the SAFE copy with IEEE divisions.  Same bits either way; the quick
             * copy is branch-free, so ptxas schedules each order as one basic block.             */
            out.line("#define DIVY_(a, b, y, ok) tb_divy_q((a), (b), (y), chk)");
            out.line("#define DIVQ_(a, b) tb_div_q((a), (b), chk)");
            out.line("#define DENCHK_(d) chk = max(chk, tb_rng(d))");
            out.line("#define TB_ORDER_SYNC() %s", (co.flags & 128) ? "__syncthreads()" : "");
            out.line("__device__ __forceinline__ unsigned tb_u_jets_q%s {", sig);
            out.line("  unsigned chk = 0u;");
            out.s += body.s;
            out.line("  return chk;");
            out.line("}");
            out.line("#undef DIVY_\n#undef DIVQ_\n#undef DENCHK_\n#undef TB_ORDER_SYNC");
        }
        out.line("#define DIVY_(a, b, y, ok) tb_div_y((a), (b), (y), (ok))");
        out.line("#define DIVQ_(a, b) D_((a), (b))");
        out.line("#define DENCHK_(d)");
        out.line("#define TB_ORDER_SYNC() %s", (!quick && (co.flags & 128)) ? "__syncthreads()" : "");
        out.line("__device__ __forceinline__ void tb_u_jets%s%s {", quick ? "_s" : "", sig);
        out.s += body.s;
        out.line("}");
        if (quick) {
            out.line("__device__ __forceinline__ void tb_u_jets%s {", sig);
            /* tensor-memory accesses are warp-collective: when any lane needs the guarded divisions the whole warp
             * takes the safe copy (same bits either way)                                                          */
            if (ntm > 0) out.line("  if (__any_sync(0xffffffffu, tb_u_jets_q(J, smt, tm, par, t0, x0) >= TB_RNG_SPAN)) tb_u_jets_s(J, smt, tm, par, t0, x0);");
            else out.line("  if (tb_u_jets_q(J, smt, tm, par, t0, x0) >= TB_RNG_SPAN) tb_u_jets_s(J, smt, tm, par, t0, x0);");
            out.line("}");
        }
    }

    /* coefficient k of state jet c; resolves to a register, an LDS or a scaled dx
     * coefficient once the caller's loops over (c, k) are unrolled                */
    void emit_accessor(Em &e) {
        const tb_ode_desc_t &d = ode.desc;
        e.line("__device__ __forceinline__ double tb_u_x(const TbJets &J, const double *__restrict__ smt, const double *__restrict__ x0, const int c, const int k) {");
        e.line("  switch (c) {");
        for (int c = 0; c < d.dof; c++) {
            const SeriesInfo &si = ser[TB_OPERAND(TB_K_X, c)];
            if (si.place == PL_REG) e.line("    case %d: return %s;", c, dynref(si, "k").c_str());
            else if (si.place == PL_SMEM && si.trim == 0) e.line("    case %d: return %s;", c, sref(si, "k").c_str());
            else if (si.place == PL_TMEM && si.trim == 0) e.line("    case %d: return tb_tm_ld(J.tm + 2 * (%d + k));", c, si.tbase);
            else if (si.place == PL_SMEM || si.place == PL_TMEM) {
                const SeriesInfo &di = ser[si.dx];
                std::string r0 = dynref(di, "0"), r1 = dynref(di, "1");
                std::string hi = si.place == PL_SMEM ? sref(si, "k") : fmt("tb_tm_ld(J.tm + 2 * (%d + k))", si.tbase);
                e.line("    case %d: return k == 0 ? x0[%d] : (k == 1 ? M_(%s, 1.0) : (k == 2 ? M_(%s, 0.5) : %s));", c, c, r0.c_str(),
                       r1.c_str(), hi.c_str());
            } else {
                const SeriesInfo &di = ser[TB_OPERAND(TB_K_DX, c)];
                std::string ref = dynref(di, "(k - 1)");
                if (fast) e.line("    case %d: return k == 0 ? x0[%d] : %s * (1.0 / (double)k);", c, c, ref.c_str());
                else e.line("    case %d: return k == 0 ? x0[%d] : tb_divc(%s, k);", c, c, ref.c_str());
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
    /* events: supported when g is a state component (Poincare-section style, e.g. g = x[2] of
     * test/rootfinding.jl:26) with an optional gate on another state component; event functions that need
     * their own tape go to variant 1                                                                      */
    const bool events = co.kind == TB_KIND_EVENTS;
    if (events) {
        const bool gate_ok = d.ev_gate_cmp == TB_GATE_ALWAYS || TB_OPERAND_KIND(d.ev_gate) == TB_K_X;
        if (d.ev_nops != 0 || TB_OPERAND_KIND(d.ev_g) != TB_K_X || !gate_ok) { err = "the unrolled generator covers events on state components"; return TB_ERR_UNSUPPORTED; }
    } else if (co.kind == TB_KIND_LYAPCOL) {
        if (d.dof < 2 || (d.dof & 1) || d.dof / 2 > 16) { err = "TB_KIND_LYAPCOL needs the column system [x; v] (an even number of components, at most 32)"; return TB_ERR_ARG; }
    } else if (co.kind != TB_KIND_PLAIN) { err = "the unrolled generator covers TB_KIND_PLAIN, TB_KIND_EVENTS and TB_KIND_LYAPCOL"; return TB_ERR_UNSUPPORTED; }
    if (co.order < 1 || co.order > 40) { err = "order must be in 1..40 for the unrolled generator"; return TB_ERR_UNSUPPORTED; }
    if ((co.flags & 128) && co.kind != TB_KIND_PLAIN) {
        /* the per-order barriers sit inside the jet body, which the event and team kernels also call from divergent code */
        err = "TB_CF_BLOCK_LOCKSTEP is available for TB_KIND_PLAIN kernels only";
        return TB_ERR_ARG;
    }
    if (d.dof > 16 || d.nops > 200) { err = "tape too large to unroll"; return TB_ERR_UNSUPPORTED; }
    for (int j = 0; j < d.nops; j++)
        if (ode.ops[j].op >= TB_OP_TAN && ode.ops[j].op <= TB_OP_ACOS) {
            err = "tan!/sinhcosh!/tanh!/atan!/asin!/acos! are lowered by the rolled generator";
            return TB_ERR_UNSUPPORTED;
        }
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
    /* defaults from the B200 sweeps of round 1: fast mode keeps 3 series in registers
     * and derives unread state jets; parity mode (more arithmetic per load, Horner on exact quotients) does
     * better with 4 series in registers and the state jets stored (TB_CF_DERIVE_X overrides)            */
    const bool parity = co.mode == TB_MODE_PARITY;
    int budget = co.reg_budget > 0 ? co.reg_budget : (parity ? 4 : 3) * np1;
    if (co.reg_budget < 0) budget = 0;
    g.derive_x = parity ? (co.flags & 512) != 0 : !(co.flags & 256);
    g.trim_x = parity && co.kind == TB_KIND_PLAIN && !(co.flags & 4096);
    g.split = (co.flags & 8192) != 0;
    g.top_trim = (co.flags >> 20) & 15; /* TB_CF_TOPTRIM(t) */
    g.pow_recip = (co.flags & TB_CF_POW_RECIP) != 0;
    g.prefetch = (co.flags >> 24) & 15; /* TB_CF_PREFETCH(d) */
    g.mulahead = parity ? (co.flags >> 28) & 7 : 0; /* TB_CF_MULAHEAD(d), lines = 2*d */
    /* default of the parity plain / team kernels (B200 sweep, round 2: Kepler-28 +5 %, Lorenz column kernel +1.6 %,
     * pendulum events +-0): loads 16 lines and products 8 lines ahead; deeper pipelines spill more than they hide */
    if (parity && g.prefetch == 0 && g.mulahead == 0 && !(co.flags & TB_CF_NO_SWPIPE) && co.kind != TB_KIND_EVENTS) { g.prefetch = 8; g.mulahead = 4; }
    g.mulahead *= 2; g.prefetch *= 2;
    /* tensor memory: 512 columns x 128 lanes of 32 bits per SM; warp w reaches lanes 32*(w%4)..+31, so the warps that
     * share a lane quadrant split the columns: 256 threads -> 256 columns = 128 doubles per thread.  TB_CF_TMEM
     * turns the tier on (plain kernels; the jet body of the event / team kernels runs in divergent code).        */
    int tmem_budget = 0;
    g.tmem_tails = (co.flags & TB_CF_TMEM_TAILS) && co.kind == TB_KIND_PLAIN && (co.flags & TB_CF_FINAL_ONLY) && parity;
    if (((co.flags & TB_CF_TMEM) || g.tmem_tails) && co.kind == TB_KIND_PLAIN) {
        const int btmax = co.block_threads > 0 ? co.block_threads : 256;
        const int share = ((btmax + 31) / 32 + 3) / 4;
        tmem_budget = (512 / share) / 2;
        if (co.tmem_budget > 0) tmem_budget = std::min(tmem_budget, co.tmem_budget);
    }
    g.analyse(budget, tmem_budget);
    L.variant = 2;
    L.nx = d.dof;
    L.nslot = g.nsm;
    L.slot.clear();
    const int evslot = g.nsm; /* the series of g gets a shared-memory slot of its own (dynamic indexing in findroot!) */
    if (events) { g.nsm += 1; g.smem_doubles += np1; } /* trim_x is off for events: the slots are whole, S(TB_EVSLOT, k) is valid */
    const size_t per_traj = (size_t)std::max(1, g.smem_doubles) * sizeof(double);
    const size_t smem_max = 227 * 1024;
    int bt = co.block_threads;
    if (bt <= 0) {
        bt = (int)(smem_max / per_traj) / 32 * 32;
        if (bt > 256) bt = 256;
    }
    if (bt < 32 || bt > 512 || (bt % 32) || (size_t)bt * per_traj > smem_max) {
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
    if (co.kind == TB_KIND_PLAIN || co.kind == TB_KIND_EVENTS) e.line("#define TB_FINAL_ONLY %d", (co.flags & 262144) ? 1 : 0);
    if (co.kind == TB_KIND_LYAPCOL) e.line("#define TB_NQ %d", d.dof / 2);
    if (co.kind == TB_KIND_LYAPCOL) e.line("#define TB_LYAP_TRANGE %d", (co.flags & 131072) ? 1 : 0);
    e.line("#define TB_EV_NOPS 0");
    e.line("#define TB_JAC_NOPS 0");
    /* opt-in only: measured -7 % on Kepler-28 (4 register series) and -3 % on pendulum-25 (3 register series) */
    e.line("#define TB_STEPSIZE_PAIR %d", (parity && (co.flags & 65536)) ? 1 : 0);
    /* step lockstep (one block barrier per step) is the default for event kernels: their step body is long and
     * branchy relative to its FP64 work, and sharing instruction fetches is worth 19 % on the pendulum workload;
     * the plain Kepler kernel gains nothing from it (+-1 %), so it stays opt-in there (TB_CF_STEP_LOCKSTEP)   */
    const bool step_lock = (co.flags & 1024) || (co.kind == TB_KIND_EVENTS && !(co.flags & 16384));
    e.line("#define TB_BAR %d", (co.flags & 128) ? 1 : step_lock ? 2 : 0);
    e.line("#define TB_TMEM %d", g.ntm > 0 ? 1 : 0);
    if (g.ntm > 0) {
        const int share = (bt / 32 + 3) / 4;
        if (2 * g.tmem_doubles > 512 / share) { err = "internal: tensor-memory window overflow"; return TB_ERR_ARG; }
        e.line("#define TB_TM_COLS %d /* columns of the per-warp window (%d warps share a lane quadrant) */", (512 / share) & ~1, share);
    }
    e.s += embedded_math_h();
    e.s += embedded_device_pre();
    e.line("/* placement: %d series in registers, %d in shared memory, %d in tensor memory */", g.nreg, g.nsm, g.ntm);
    e.line("struct TbJets {");
    for (int i = 0; i < g.nreg; i++) {
        int kr = np1;
        for (auto &kv : g.ser) if (kv.second.place == PL_REG && kv.second.id == i) kr = kv.second.kreg;
        e.line("  double r%d[%d];", i, kr);
    }
    e.line("  unsigned tm; /* this warp's tensor-memory window (TB_TMEM kernels) */");
    e.line("};");
    g.emit_jets(e);
    g.emit_accessor(e);
    if (events) {
        const int cg = TB_OPERAND_INDEX(d.ev_g);
        e.line("#define TB_EVSLOT %d", evslot);
        e.line("#define TB_EV_G(kk) S(TB_EVSLOT, (kk))");
        const char *cmp[] = {"", ">", "<", ">=", "<="};
        if (d.ev_gate_cmp == TB_GATE_ALWAYS) e.line("#define TB_EV_GATE_EXPR true");
        else e.line("#define TB_EV_GATE_EXPR (tb_u_x(J, smt, x0, %d, 0) %s 0.0)", TB_OPERAND_INDEX(d.ev_gate), cmp[d.ev_gate_cmp]);
        /* order-0 forms on the state alone: what tb_traj_init needs when eventorder == 0 (no jets at refill) */
        e.line("#define TB_EV_G0(x0) ((x0)[%d])", cg);
        if (d.ev_gate_cmp == TB_GATE_ALWAYS) e.line("#define TB_EV_GATE0(x0) true");
        else e.line("#define TB_EV_GATE0(x0) ((x0)[%d] %s 0.0)", TB_OPERAND_INDEX(d.ev_gate), cmp[d.ev_gate_cmp]);
        e.line("/* coefficient k (run-time) of g = x[%d] */", cg);
        e.line("__device__ __forceinline__ double tb_u_gcoef(const TbJets &J, const double *__restrict__ smt, const double *__restrict__ x0, const int k) {");
        e.line("  switch (k) {");
        for (int k = 0; k <= co.order; k++) e.line("    case %d: return tb_u_x(J, smt, x0, %d, %d);", k, cg, k);
        e.line("    default: return 0.0;");
        e.line("  }");
        e.line("}");
        e.line("__device__ __forceinline__ void tb_u_gcopy(const TbJets &J, double *__restrict__ smt, const double *__restrict__ x0) {");
        for (int k = 0; k <= co.order; k++) e.line("  S(TB_EVSLOT, %d) = tb_u_x(J, smt, x0, %d, %d);", k, cg, k);
        e.line("}");
    }
    e.s += embedded_device_drv();
    if (!g.err.empty()) { err = g.err; return TB_ERR_ARG; }
    src.swap(e.s);
    return TB_OK;
}

} // namespace tb
