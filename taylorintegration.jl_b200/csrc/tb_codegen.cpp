/*
 * tb_codegen.cpp -- tape -> CUDA C++ (the B200 counterpart of what `@taylorize`
 * does in the reference: src/parse_eqs.jl:236-358 turns an RHS into a sequence
 * of TaylorSeries.<op>!(res, a, b, ord) calls inside `for ord = 0:order-1`;
 * here the same sequence is lowered to straight-line device code that NVRTC
 * fuses with the driver loop into one kernel per ODE).
 *
 * Storage analysis: a tape value needs a whole series in shared memory only if
 * some recurrence reads it at an index other than the current order (operands
 * of Cauchy products, results of div!/pow!/sqrt!/exp!/log!/sincos!), or if a
 * later segment (event function, Jacobian) reads it after the jets are
 * complete.  Everything else -- the reference allocates a full Taylor1 for each
 * of them (RetAlloc, parse_eqs.jl:80-107) -- is a register scalar.
 */
#include "tb_internal.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <set>
#include <sstream>

namespace tb {

namespace {

struct Emitter {
    std::string s;
    void line(const char *fmt, ...) {
        char buf[2048];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        s += buf;
        s += "\n";
    }
};

std::string lit(double v) {
    unsigned long long u;
    memcpy(&u, &v, 8);
    char buf[96];
    snprintf(buf, sizeof buf, "__longlong_as_double(0x%016llxLL)/*%.17g*/", u, v);
    return buf;
}

bool is_num(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_PAR || k == TB_K_IMM; }
bool is_series(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_X || k == TB_K_DX || k == TB_K_TMP || k == TB_K_TIME; }

/* operands an op reads at indices other than the current order (incl. its own result) */
void index_reads(const tb_op_t &op, std::vector<int32_t> &out) {
    switch (op.op) {
        case TB_OP_MUL: if (is_series(op.a) && is_series(op.b)) { out.push_back(op.a); out.push_back(op.b); } break;
        case TB_OP_SQR: out.push_back(op.a); break;
        case TB_OP_DIV: if (is_series(op.b)) { out.push_back(op.b); out.push_back(op.res); } break;
        case TB_OP_SQRT: out.push_back(op.res); break;
        case TB_OP_POW:
            if (op.imm == 2.0) out.push_back(op.a);
            else if (op.imm == 0.5) out.push_back(op.res);
            else if (op.imm != 0.0 && op.imm != 1.0) { out.push_back(op.a); out.push_back(op.res); }
            break;
        case TB_OP_EXP: case TB_OP_LOG: out.push_back(op.a); out.push_back(op.res); break;
        case TB_OP_SINCOS: case TB_OP_SINHCOSH: out.push_back(op.a); out.push_back(op.res); out.push_back(op.aux); break;
        case TB_OP_TAN: case TB_OP_TANH: case TB_OP_ATAN: case TB_OP_ASIN: case TB_OP_ACOS:
            out.push_back(op.a); out.push_back(op.res); out.push_back(op.aux); break;
        default: break;
    }
}

struct Gen {
    const Ode &ode;
    const tb_compile_opts_t &co;
    Layout &L;
    int P;
    std::string err;

    Gen(const Ode &o, const tb_compile_opts_t &c, Layout &l) : ode(o), co(c), L(l), P(c.order) {}

    int slot_of(int32_t o) const {
        if (TB_OPERAND_KIND(o) == TB_K_X) return TB_OPERAND_INDEX(o);
        auto it = L.slot.find(o);
        return it == L.slot.end() ? -1 : it->second;
    }
    void mark(int32_t o) {
        if (!is_series(o) || TB_OPERAND_KIND(o) == TB_K_X) return;
        if (L.slot.find(o) == L.slot.end()) L.slot[o] = L.nslot++;
    }

    void analyse() {
        const tb_ode_desc_t &d = ode.desc;
        L.nx = (co.kind == TB_KIND_LYAP) ? d.dof + d.dof * d.dof : d.dof;
        L.nslot = L.nx;
        L.slot.clear();
        const int seg_first[3] = {0, d.nops, d.nops + d.ev_nops};
        const int seg_count[3] = {d.nops, co.kind == TB_KIND_EVENTS ? d.ev_nops : 0, co.kind == TB_KIND_LYAP ? d.jac_nops : 0};
        for (int sgm = 0; sgm < 3; sgm++) {
            std::set<int32_t> produced;
            for (int j = seg_first[sgm]; j < seg_first[sgm] + seg_count[sgm]; j++) {
                const tb_op_t &op = ode.ops[j];
                std::vector<int32_t> rd;
                index_reads(op, rd);
                for (int32_t o : rd) mark(o);
                if (sgm > 0) { /* later segments run after the jets are complete */
                    const int32_t srcs[2] = {op.a, op.b};
                    for (int32_t o : srcs)
                        if (is_series(o) && !produced.count(o)) mark(o);
                }
                produced.insert(op.res);
                if (op.op == TB_OP_SINCOS || (op.op >= TB_OP_TAN && op.op <= TB_OP_ACOS)) produced.insert(op.aux);
            }
        }
        if (co.kind == TB_KIND_EVENTS) {
            mark(d.ev_g);
            if (d.ev_gate_cmp != TB_GATE_ALWAYS) mark(d.ev_gate);
        }
        if (co.kind == TB_KIND_LYAP)
            for (int i = 0; i < d.dof * d.dof; i++)
                if (ode.jac[i].kind == TB_JAC_SERIES) mark(ode.jac[i].operand);
    }

    static std::string vname(int32_t o) {
        char b[32];
        int k = TB_OPERAND_KIND(o), i = TB_OPERAND_INDEX(o);
        snprintf(b, sizeof b, "%s%d", k == TB_K_DX ? "vD" : (k == TB_K_TMP ? "vT" : (k == TB_K_X ? "vX" : "vTIME")), i);
        return b;
    }
    std::string num(int32_t o, const tb_op_t &op) const {
        if (TB_OPERAND_KIND(o) == TB_K_PAR) { char b[32]; snprintf(b, sizeof b, "par[%d]", TB_OPERAND_INDEX(o)); return b; }
        return lit(op.imm);
    }
    /* stored series at an arbitrary index expression */
    std::string at(int32_t o, const std::string &idx) {
        int s = slot_of(o);
        if (s < 0) { err = "internal: operand " + vname(o) + " not stored (index " + idx + ")"; return "0.0"; }
        char b[128];
        snprintf(b, sizeof b, "S(%d, %s)", s, idx.c_str());
        return b;
    }
    /* series operand at the current order */
    std::string cur(int32_t o, bool k0, const std::set<int32_t> &defined) {
        int kind = TB_OPERAND_KIND(o);
        if (kind == TB_K_TIME && slot_of(o) < 0) return k0 ? "t0" : "(k == 1 ? 1.0 : 0.0)";
        if (kind != TB_K_X && defined.count(o)) return vname(o);
        if (slot_of(o) >= 0) return at(o, k0 ? "0" : "k");
        err = "tape reads " + vname(o) + " before it is defined";
        return "0.0";
    }

    void emit_op(Emitter &e, const tb_op_t &op, bool k0, std::set<int32_t> &defined) {
        const bool sa = is_num(op.a), sb = is_num(op.b);
        const std::string r = vname(op.res);
        std::string a_k, b_k;
        if (TB_OPERAND_KIND(op.a) != TB_K_NONE) a_k = sa ? num(op.a, op) : cur(op.a, k0, defined);
        if (TB_OPERAND_KIND(op.b) != TB_K_NONE) b_k = sb ? num(op.b, op) : cur(op.b, k0, defined);
        const char *A = a_k.c_str(), *B = b_k.c_str(), *R = r.c_str();
        bool stored_early = false; /* result must be in smem before the recurrence reads it? (never: recurrences read < k) */
        (void)stored_early;
        int opc = op.op;
        double pw = op.imm;
        if (opc == TB_OP_POW) {
            if (pw == 0.0) opc = TB_OP_ONE;
            else if (pw == 1.0) opc = TB_OP_IDENTITY;
            else if (pw == 2.0) opc = TB_OP_SQR;
            else if (pw == 0.5) opc = TB_OP_SQRT;
        }
        switch (opc) {
            case TB_OP_IDENTITY:
                if (sa) e.line("  %s = %s;", R, k0 ? A : "0.0");
                else e.line("  %s = %s;", R, A);
                break;
            case TB_OP_NEG: e.line("  %s = N_(%s);", R, A); break;
            case TB_OP_ZERO: e.line("  %s = 0.0;", R); break;
            case TB_OP_ONE: e.line("  %s = %s;", R, k0 ? "1.0" : "0.0"); break;
            case TB_OP_ADD:
            case TB_OP_SUB: {
                const char *F = opc == TB_OP_ADD ? "A_" : "B_";
                if (sa && sb) { if (k0) e.line("  %s = %s(%s, %s);", R, F, A, B); else e.line("  %s = 0.0;", R); }
                else if (sa) {
                    if (k0) e.line("  %s = %s(%s, %s);", R, F, A, B);
                    else if (opc == TB_OP_ADD) e.line("  %s = %s;", R, B);
                    else e.line("  %s = N_(%s);", R, B);
                } else if (sb) { if (k0) e.line("  %s = %s(%s, %s);", R, F, A, B); else e.line("  %s = %s;", R, A); }
                else e.line("  %s = %s(%s, %s);", R, F, A, B);
                break;
            }
            case TB_OP_MUL:
                if (sa && sb) { if (k0) e.line("  %s = M_(%s, %s);", R, A, B); else e.line("  %s = 0.0;", R); }
                else if (sa || sb) e.line("  %s = M_(%s, %s);", R, A, B);
                else if (k0) e.line("  %s = M_(%s, %s);", R, A, B);
                else {
                    e.line("  { double c = M_(%s, %s);", at(op.a, "0").c_str(), B);
                    e.line("    for (int i = 1; i <= k; i++) c = A_(c, M_(%s, %s));", at(op.a, "i").c_str(), at(op.b, "k - i").c_str());
                    e.line("    %s = c; }", R);
                }
                break;
            case TB_OP_SQR:
                if (k0) e.line("  %s = M_(%s, %s);", R, A, A);
                else {
                    e.line("  { const int kodd = k & 1, kend = (k - 2 + kodd) / 2;");
                    e.line("    double c = M_(%s, %s);", at(op.a, "0").c_str(), A);
                    e.line("    for (int i = 1; i <= kend; i++) c = A_(c, M_(%s, %s));", at(op.a, "i").c_str(), at(op.a, "k - i").c_str());
                    e.line("    c = M_(2.0, c);");
                    e.line("    if (!kodd) { const double h = %s; c = A_(c, M_(h, h)); }", at(op.a, "k >> 1").c_str());
                    e.line("    %s = c; }", R);
                }
                break;
            case TB_OP_DIV:
                if (sa && sb) { if (k0) e.line("  %s = D_(%s, %s);", R, A, B); else e.line("  %s = 0.0;", R); }
                else if (sb) e.line("  %s = D_(%s, %s);", R, A, B);
                else if (k0) e.line("  %s = D_(%s, %s);", R, A, B);
                else {
                    e.line("  { double s = M_(%s, %s);", at(op.res, "0").c_str(), B);
                    e.line("    for (int i = 1; i <= k - 1; i++) s = A_(s, M_(%s, %s));", at(op.res, "i").c_str(), at(op.b, "k - i").c_str());
                    if (sa) e.line("    %s = D_(N_(s), %s); }", R, at(op.b, "0").c_str());
                    else e.line("    %s = D_(B_(%s, s), %s); }", R, A, at(op.b, "0").c_str());
                }
                break;
            case TB_OP_SQRT:
                if (k0) e.line("  %s = Q_(%s);", R, A);
                else {
                    e.line("  { const int kodd = k & 1, kend = (k - 2 + kodd) / 2;");
                    e.line("    double s = 0.0;");
                    e.line("    if (kend >= 1) s = M_(%s, %s);", at(op.res, "1").c_str(), at(op.res, "k - 1").c_str());
                    e.line("    for (int i = 2; i <= kend; i++) s = A_(s, M_(%s, %s));", at(op.res, "i").c_str(), at(op.res, "k - i").c_str());
                    e.line("    double aux = B_(%s, M_(2.0, s));", A);
                    e.line("    if (!kodd) { const double h = %s; aux = B_(aux, M_(h, h)); }", at(op.res, "kend + 1").c_str());
                    e.line("    %s = D_(aux, M_(2.0, %s)); }", R, at(op.res, "0").c_str());
                }
                break;
            case TB_OP_POW:
                if (k0) e.line("  %s = tb_pow(%s, %s);", R, A, lit(pw).c_str());
                else {
                    e.line("  { const double pw = %s;", lit(pw).c_str());
                    e.line("    double s = M_(M_(M_(pw, (double)k), %s), %s);", at(op.res, "0").c_str(), A);
                    e.line("    for (int i = 1; i <= k - 1; i++) { const double rr = B_(M_(pw, (double)(k - i)), (double)i);");
                    e.line("      s = A_(s, M_(M_(rr, %s), %s)); }", at(op.res, "i").c_str(), at(op.a, "k - i").c_str());
                    e.line("    %s = D_(s, M_((double)k, %s)); }", R, at(op.a, "0").c_str());
                }
                break;
            case TB_OP_EXP:
                if (k0) e.line("  %s = tb_exp(%s);", R, A);
                else {
                    e.line("  { double s = M_(M_((double)k, %s), %s);", A, at(op.res, "0").c_str());
                    e.line("    for (int i = 1; i <= k - 1; i++) s = A_(s, M_(M_((double)(k - i), %s), %s));", at(op.a, "k - i").c_str(), at(op.res, "i").c_str());
                    e.line("    %s = D_(s, (double)k); }", R);
                }
                break;
            case TB_OP_LOG:
                if (k0) e.line("  %s = tb_log(%s);", R, A);
                else {
                    e.line("  if (k == 1) %s = D_(%s, %s);", R, A, at(op.a, "0").c_str());
                    e.line("  else { double s = M_(M_((double)(k - 1), %s), %s);", at(op.a, "1").c_str(), at(op.res, "k - 1").c_str());
                    e.line("    for (int i = 2; i <= k - 1; i++) s = A_(s, M_(M_((double)(k - i), %s), %s));", at(op.a, "i").c_str(), at(op.res, "k - i").c_str());
                    e.line("    %s = D_(B_(%s, D_(s, (double)k)), %s); }", R, A, at(op.a, "0").c_str());
                }
                break;
            case TB_OP_SINCOS: {
                const std::string rc = vname(op.aux);
                if (k0) e.line("  tb_sincos(%s, &%s, &%s);", A, R, rc.c_str());
                else {
                    e.line("  { double x = %s;", at(op.a, "1").c_str());
                    e.line("    double sk = M_(x, %s), ck = M_(N_(x), %s);", at(op.aux, "k - 1").c_str(), at(op.res, "k - 1").c_str());
                    e.line("    for (int i = 2; i <= k; i++) { x = M_((double)i, %s);", at(op.a, "i").c_str());
                    e.line("      sk = A_(sk, M_(x, %s)); ck = B_(ck, M_(x, %s)); }", at(op.aux, "k - i").c_str(), at(op.res, "k - i").c_str());
                    e.line("    %s = D_(sk, (double)k); %s = D_(ck, (double)k); }", R, rc.c_str());
                }
                if (slot_of(op.aux) >= 0) e.line("  %s = %s;", at(op.aux, k0 ? "0" : "k").c_str(), rc.c_str());
                defined.insert(op.aux);
                break;
            }
            case TB_OP_SINHCOSH: { /* sinhcosh!(s, c, a, k) */
                const std::string rc = vname(op.aux);
                if (k0) e.line("  tb_sinhcosh(%s, &%s, &%s);", A, R, rc.c_str());
                else {
                    e.line("  { double x = %s;", at(op.a, "1").c_str());
                    e.line("    double sk = M_(x, %s), ck = M_(x, %s);", at(op.aux, "k - 1").c_str(), at(op.res, "k - 1").c_str());
                    e.line("    for (int i = 2; i <= k; i++) { x = M_((double)i, %s);", at(op.a, "i").c_str());
                    e.line("      sk = A_(sk, M_(x, %s)); ck = A_(ck, M_(x, %s)); }", at(op.aux, "k - i").c_str(), at(op.res, "k - i").c_str());
                    e.line("    %s = D_(sk, (double)k); %s = D_(ck, (double)k); }", R, rc.c_str());
                }
                e.line("  %s = %s;", at(op.aux, k0 ? "0" : "k").c_str(), rc.c_str());
                defined.insert(op.aux);
                break;
            }
            case TB_OP_TAN: case TB_OP_TANH: { /* tan!(c, a, c2, k), tanh!(c, a, c2, k): aux = c^2 */
                const std::string rc = vname(op.aux);
                const bool hyp = opc == TB_OP_TANH;
                if (k0) { e.line("  %s = %s(%s);", R, hyp ? "tb_tanh" : "tb_tan", A); e.line("  %s = M_(%s, %s);", rc.c_str(), R, R); }
                else {
                    e.line("  { double s = M_(M_((double)k, %s), %s);", A, at(op.aux, "0").c_str());
                    e.line("    for (int i = 1; i <= k - 1; i++) s = A_(s, M_(M_((double)(k - i), %s), %s));", at(op.a, "k - i").c_str(), at(op.aux, "i").c_str());
                    e.line("    %s = %s(%s, D_(s, (double)k)); }", R, hyp ? "B_" : "A_", A);
                    /* sqr!(c2, c, k) with c[k] = the value just computed */
                    e.line("  { const int kodd = k & 1, kend = (k - 2 + kodd) / 2;");
                    e.line("    double c = M_(%s, %s);", at(op.res, "0").c_str(), R);
                    e.line("    for (int i = 1; i <= kend; i++) c = A_(c, M_(%s, %s));", at(op.res, "i").c_str(), at(op.res, "k - i").c_str());
                    e.line("    c = M_(2.0, c);");
                    e.line("    if (!kodd) { const double h = %s; c = A_(c, M_(h, h)); }", at(op.res, "k >> 1").c_str());
                    e.line("    %s = c; }", rc.c_str());
                }
                e.line("  %s = %s;", at(op.aux, k0 ? "0" : "k").c_str(), rc.c_str());
                defined.insert(op.aux);
                break;
            }
            case TB_OP_ATAN: case TB_OP_ASIN: case TB_OP_ACOS: { /* aux r = 1 + a^2 (atan!) or sqrt(1 - a^2) (asin!, acos!) */
                const std::string rc = vname(op.aux);
                if (k0) {
                    e.line("  %s = %s(%s);", R, opc == TB_OP_ATAN ? "tb_atan" : (opc == TB_OP_ASIN ? "tb_asin" : "tb_acos"), A);
                    if (opc == TB_OP_ATAN) e.line("  %s = A_(1.0, M_(%s, %s));", rc.c_str(), A, A);
                    else e.line("  %s = Q_(B_(1.0, M_(%s, %s)));", rc.c_str(), A, A);
                } else {
                    e.line("  { double s = 0.0;");
                    e.line("    if (k >= 2) s = M_(M_((double)(k - 1), %s), %s);", at(op.aux, "1").c_str(), at(op.res, "k - 1").c_str());
                    e.line("    for (int i = 2; i <= k - 1; i++) s = A_(s, M_(M_((double)(k - i), %s), %s));", at(op.aux, "i").c_str(), at(op.res, "k - i").c_str());
                    /* (a^2)[k] by sqr! */
                    e.line("    const int kodd = k & 1, kend = (k - 2 + kodd) / 2;");
                    e.line("    double c = M_(%s, %s);", at(op.a, "0").c_str(), A);
                    e.line("    for (int i = 1; i <= kend; i++) c = A_(c, M_(%s, %s));", at(op.a, "i").c_str(), at(op.a, "k - i").c_str());
                    e.line("    c = M_(2.0, c);");
                    e.line("    if (!kodd) { const double h = %s; c = A_(c, M_(h, h)); }", at(op.a, "k >> 1").c_str());
                    if (opc == TB_OP_ATAN) e.line("    %s = c;", rc.c_str());
                    else { /* sqrt!(r, 1 - a^2, k) */
                        e.line("    double q = 0.0;");
                        e.line("    if (kend >= 1) q = M_(%s, %s);", at(op.aux, "1").c_str(), at(op.aux, "k - 1").c_str());
                        e.line("    for (int i = 2; i <= kend; i++) q = A_(q, M_(%s, %s));", at(op.aux, "i").c_str(), at(op.aux, "k - i").c_str());
                        e.line("    double aux = B_(N_(c), M_(2.0, q));");
                        e.line("    if (!kodd) { const double h = %s; aux = B_(aux, M_(h, h)); }", at(op.aux, "kend + 1").c_str());
                        e.line("    %s = D_(aux, M_(2.0, %s));", rc.c_str(), at(op.aux, "0").c_str());
                    }
                    if (opc == TB_OP_ACOS) e.line("    %s = D_(N_(A_(%s, D_(s, (double)k))), %s); }", R, A, at(op.aux, "0").c_str());
                    else e.line("    %s = D_(B_(%s, D_(s, (double)k)), %s); }", R, A, at(op.aux, "0").c_str());
                }
                e.line("  %s = %s;", at(op.aux, k0 ? "0" : "k").c_str(), rc.c_str());
                defined.insert(op.aux);
                break;
            }
            default: err = "opcode not supported by the code generator"; break;
        }
        if (slot_of(op.res) >= 0) e.line("  %s = %s;", at(op.res, k0 ? "0" : "k").c_str(), R);
        defined.insert(op.res);
    }

    /* one function per (segment, order class) */
    void emit_segment(Emitter &e, const char *name, int first, int count, bool k0, bool solcoeff) {
        const tb_ode_desc_t &d = ode.desc;
        if (k0) e.line("__device__ __forceinline__ void %s_k0(double *__restrict__ smt, const double *__restrict__ par, const double t0) {", name);
        else e.line("__device__ __forceinline__ void %s_k(double *__restrict__ smt, const double *__restrict__ par, const double t0, const int k) {", name);
        std::set<int32_t> vars;
        for (int j = first; j < first + count; j++) {
            vars.insert(ode.ops[j].res);
            if (ode.ops[j].op == TB_OP_SINCOS || (ode.ops[j].op >= TB_OP_TAN && ode.ops[j].op <= TB_OP_ACOS)) vars.insert(ode.ops[j].aux);
        }
        for (int32_t v : vars) e.line("  double %s;", vname(v).c_str());
        std::set<int32_t> defined;
        if (solcoeff) {
            auto it = L.slot.find(TB_OPERAND(TB_K_TIME, 0));
            if (it != L.slot.end()) { /* t = t0 + Taylor1(order): init_expansions, cache.jl:176-183 */
                if (k0) {
                    e.line("  S(%d, 0) = t0;", it->second);
                    if (P >= 1) e.line("  S(%d, 1) = 1.0;", it->second);
                    e.line("  for (int j = 2; j <= TB_P; j++) S(%d, j) = 0.0;", it->second);
                }
            }
        }
        for (int j = first; j < first + count; j++) emit_op(e, ode.ops[j], k0, defined);
        if (solcoeff) { /* solcoeff!: src/TaylorIntegration.jl:33-40   x[k+1] = dx[k]/(k+1) */
            for (int c = 0; c < d.dof; c++) {
                int32_t dx = TB_OPERAND(TB_K_DX, c);
                if (!defined.count(dx)) { err = "tape does not define dx[" + std::to_string(c) + "]"; continue; }
                if (k0) e.line("  S(%d, 1) = D_(%s, 1.0);", c, vname(dx).c_str());
                else e.line("  S(%d, k + 1) = D_(%s, (double)(k + 1));", c, vname(dx).c_str());
            }
        }
        e.line("}");
    }

    /* lyap_jetcoeffs!: src/lyapunovspectrum.jl:161-211, loop order i3 outer, i2, i1 inner */
    void emit_var(Emitter &e, bool k0) {
        const int dof = ode.desc.dof;
        if (k0) e.line("__device__ __forceinline__ void tb_var_k0(double *__restrict__ smt, const double *__restrict__ par) {");
        else e.line("__device__ __forceinline__ void tb_var_k(double *__restrict__ smt, const double *__restrict__ par, const int k) {");
        for (int i = 0; i < dof * dof; i++) e.line("  double d%d = 0.0;", i);
        for (int i3 = 0; i3 < dof; i3++)
            for (int i2 = 0; i2 < dof; i2++)
                for (int i1 = 0; i1 < dof; i1++) {
                    const tb_jac_entry_t &J = ode.jac[i1 * dof + i3];
                    const int phi = dof + i3 + i2 * dof, dd = i1 + i2 * dof;
                    char ps[64];
                    if (J.kind == TB_JAC_ZERO) continue;
                    if (J.kind == TB_JAC_CONST || J.kind == TB_JAC_PAR) {
                        snprintf(ps, sizeof ps, "S(%d, %s)", phi, k0 ? "0" : "k");
                        std::string cv = lit(J.c);
                        if (J.kind == TB_JAC_PAR) {
                            char pb[160];
                            snprintf(pb, sizeof pb, "M_(%s, par[%d])", lit(J.c).c_str(), TB_OPERAND_INDEX(J.operand));
                            cv = pb;
                        }
                        e.line("  d%d = A_(d%d, M_(%s, %s));", dd, dd, cv.c_str(), ps);
                    } else if (k0) {
                        e.line("  d%d = A_(d%d, M_(%s, S(%d, 0)));", dd, dd, at(J.operand, "0").c_str(), phi);
                    } else {
                        e.line("  { double c = M_(%s, S(%d, k));", at(J.operand, "0").c_str(), phi);
                        e.line("    for (int i = 1; i <= k; i++) c = A_(c, M_(%s, S(%d, k - i)));", at(J.operand, "i").c_str(), phi);
                        e.line("    d%d = A_(d%d, c); }", dd, dd);
                    }
                }
        for (int i = 0; i < dof * dof; i++) {
            if (k0) e.line("  S(%d, 1) = d%d;", dof + i, i);
            else e.line("  S(%d, k + 1) = D_(d%d, (double)(k + 1));", dof + i, i);
        }
        e.line("}");
    }
};

} // namespace

int generate(const Ode &ode, const tb_compile_opts_t &co, Layout &L, std::string &src, std::string &err) {
    const tb_ode_desc_t &d = ode.desc;
    if (co.order < 1 || co.order > 64) { err = "order must be in 1..64"; return TB_ERR_ARG; }
    if (co.kind == TB_KIND_LYAPCOL) { err = "TB_KIND_LYAPCOL is built by the register-tiled generator only (tape too large, or TB_CF_ROLLED set)"; return TB_ERR_UNSUPPORTED; }
    if (co.kind == TB_KIND_LYAP && !d.has_jac) { err = "TB_KIND_LYAP needs a Jacobian description"; return TB_ERR_ARG; }
    if (co.kind == TB_KIND_EVENTS && TB_OPERAND_KIND(d.ev_g) == TB_K_NONE) { err = "TB_KIND_EVENTS needs an event function"; return TB_ERR_ARG; }
    Gen g(ode, co, L);
    L.variant = 1;
    g.analyse();
    /* block size: as many trajectories per SM as shared memory holds */
    const size_t per_traj = (size_t)L.nslot * (co.order + 1) * sizeof(double);
    const size_t smem_max = 227 * 1024;
    int bt = co.block_threads;
    if (bt <= 0) {
        bt = (int)(smem_max / per_traj) / 32 * 32;
        if (bt > 256) bt = 256;
    }
    if (bt < 32 || (bt % 32) || (size_t)bt * per_traj > smem_max) {
        err = "working set of " + std::to_string(per_traj) + " B per trajectory does not fit one warp in shared memory";
        return TB_ERR_UNSUPPORTED;
    }
    L.block_threads = bt;
    L.smem_bytes = (size_t)bt * per_traj;
    L.blocks_per_sm = (int)(smem_max / L.smem_bytes);
    if (L.blocks_per_sm < 1) L.blocks_per_sm = 1;
    if (co.min_blocks_per_sm > 0 && co.min_blocks_per_sm < L.blocks_per_sm) L.blocks_per_sm = co.min_blocks_per_sm;
    if (L.blocks_per_sm * bt > 1024) L.blocks_per_sm = 1024 / bt > 0 ? 1024 / bt : 1;

    Emitter e;
    e.line("/* generated by libtaylorb200 tb_codegen.cpp -- do not edit */");
    e.line("#define TB_VARIANT 1");
    e.line("#define TB_P %d", co.order);
    e.line("#define TB_DOF %d", d.dof);
    e.line("#define TB_NX %d", L.nx);
    e.line("#define TB_NPAR %d", d.npar);
    e.line("#define TB_NSLOT %d", L.nslot);
    e.line("#define TB_BT %d", bt);
    e.line("#define TB_MINB %d", L.blocks_per_sm);
    e.line("#define TB_MODE %d", co.mode);
    e.line("#define TB_KIND %d", co.kind);
    e.line("#define TB_EV_NOPS %d", co.kind == TB_KIND_EVENTS ? d.ev_nops : 0);
    e.line("#define TB_JAC_NOPS %d", co.kind == TB_KIND_LYAP ? d.jac_nops : 0);
    if (co.kind == TB_KIND_EVENTS) {
        e.line("#define TB_EV_G(kk) S(%d, (kk))", g.slot_of(d.ev_g));
        const char *cmp[] = {"", ">", "<", ">=", "<="};
        if (d.ev_gate_cmp == TB_GATE_ALWAYS) e.line("#define TB_EV_GATE_EXPR true");
        else e.line("#define TB_EV_GATE_EXPR (S(%d, 0) %s 0.0)", g.slot_of(d.ev_gate), cmp[d.ev_gate_cmp]);
    }
    e.s += embedded_math_h();
    e.s += embedded_device_pre();
    g.emit_segment(e, "tb_rhs", 0, d.nops, true, true);
    g.emit_segment(e, "tb_rhs", 0, d.nops, false, true);
    if (co.kind == TB_KIND_EVENTS && d.ev_nops > 0) {
        g.emit_segment(e, "tb_ev", d.nops, d.ev_nops, true, false);
        g.emit_segment(e, "tb_ev", d.nops, d.ev_nops, false, false);
    }
    if (co.kind == TB_KIND_LYAP) {
        if (d.jac_nops > 0) {
            g.emit_segment(e, "tb_jac", d.nops + d.ev_nops, d.jac_nops, true, false);
            g.emit_segment(e, "tb_jac", d.nops + d.ev_nops, d.jac_nops, false, false);
        }
        g.emit_var(e, true);
        g.emit_var(e, false);
    }
    e.s += embedded_device_drv();
    if (!g.err.empty()) { err = g.err; return TB_ERR_ARG; }
    src.swap(e.s);
    return TB_OK;
}

int generate_any(const Ode &ode, const tb_compile_opts_t &co, Layout &L, std::string &src, std::string &err) {
    if (!(co.flags & TB_CF_ROLLED)) {
        std::string e2;
        Layout L2;
        int rc = generate_unrolled(ode, co, L2, src, e2);
        if (rc == TB_OK) { L = L2; return rc; }
        if (rc != TB_ERR_UNSUPPORTED) { err = e2; return rc; }
    }
    return generate(ode, co, L, src, err);
}

} // namespace tb
