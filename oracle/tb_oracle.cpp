/*
 * tb_oracle.cpp -- CPU restatement of the TaylorIntegration.jl hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (libtaylorb200.so, the
 * python package) links, imports or calls this file; it is used by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * as the checker and as the timed CPU baseline.
 *
 * The reference is pure Julia and cannot run here (no julia binary, and its
 * arithmetic core TaylorSeries.jl 0.22.x is not vendored under
 * /root/reference), so this is a restatement, not a build of the reference:
 *   - driver loops, step-size control, root finding, Lyapunov spectrum follow
 *     the reference sources line by line (citations on every function, paths
 *     relative to the reference tree);
 *   - the Taylor1 recurrences (mul!, sqr!, div!, pow!, ...) restate the
 *     published TaylorSeries.jl algorithms (external dependency, compat
 *     0.22.1 in Project.toml:48): left-to-right sums, separate multiply and
 *     add, true divisions.
 * Parity status: PARTIALLY PINNED.  Pinned by the reference's own goldens
 * (tests/test_oracle_kat.py): exact 3^(k+1) jet coefficients, step counts
 * 14/28, Lorenz exponents at t=20, pendulum event times k*T/2, closed-form
 * solutions.  Not pinned by any in-tree reference test: the summation order
 * inside each primitive and Julia's own pow/sincos/log bits (we use the
 * deterministic functions of tb_math.h, < 0.6 ulp from the exact value).
 *
 * Build: g++ -O2 -ffp-contract=off -shared -fPIC (see oracle/Makefile).
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include "../include/taylorb200.h"
#include "../taylorintegration.jl_b200/csrc/tb_math.h"

namespace {

enum { MATH_TB = 0, MATH_LIBM = 1 };

struct Math {
    int mode;
    double pow_(double x, double y) const { return mode == MATH_TB ? tb_pow(x, y) : std::pow(x, y); }
    double exp_(double x) const { return mode == MATH_TB ? tb_exp(x) : std::exp(x); }
    double log_(double x) const { return mode == MATH_TB ? tb_log(x) : std::log(x); }
    double tan_(double x) const { return mode == MATH_TB ? tb_tan(x) : std::tan(x); }
    double tanh_(double x) const { return mode == MATH_TB ? tb_tanh(x) : std::tanh(x); }
    double atan_(double x) const { return mode == MATH_TB ? tb_atan(x) : std::atan(x); }
    double asin_(double x) const { return mode == MATH_TB ? tb_asin(x) : std::asin(x); }
    double acos_(double x) const { return mode == MATH_TB ? tb_acos(x) : std::acos(x); }
    void sinhcosh_(double x, double *s, double *c) const {
        if (mode == MATH_TB) tb_sinhcosh(x, s, c);
        else { *s = std::sinh(x); *c = std::cosh(x); }
    }
    void sincos_(double x, double *s, double *c) const {
        if (mode == MATH_TB) tb_sincos(x, s, c);
        else { *s = std::sin(x); *c = std::cos(x); }
    }
};

/* Working set of one trajectory: the reference's VectorCache
 * (src/integrator/cache.jl:33-60): t, x, dx jets plus the RetAlloc
 * temporaries (src/parse_eqs.jl:80-107).                                     */
struct Work {
    const tb_ode_desc_t *d;
    int P;            /* order */
    int np1;          /* P + 1 */
    Math math;
    std::vector<double> t, x, dx, tmp, par;
    double imm_scratch;

    Work(const tb_ode_desc_t *desc, int order, int math_mode, int nstate_extra = 0)
        : d(desc), P(order), np1(order + 1) {
        math.mode = math_mode;
        t.assign(np1, 0.0);
        x.assign((size_t)(d->dof + nstate_extra) * np1, 0.0);
        dx.assign((size_t)(d->dof + nstate_extra) * np1, 0.0);
        tmp.assign((size_t)std::max(1, d->ntmp) * np1, 0.0);
        par.assign(std::max(1, d->npar), 0.0);
    }
    double *X(int i) { return &x[(size_t)i * np1]; }
    double *DX(int i) { return &dx[(size_t)i * np1]; }
    double *series(int32_t o) {
        int idx = TB_OPERAND_INDEX(o);
        switch (TB_OPERAND_KIND(o)) {
            case TB_K_X: return &x[(size_t)idx * np1];
            case TB_K_DX: return &dx[(size_t)idx * np1];
            case TB_K_TMP: return &tmp[(size_t)idx * np1];
            case TB_K_TIME: return t.data();
            default: return nullptr;
        }
    }
    bool is_scalar(int32_t o) const { int k = TB_OPERAND_KIND(o); return k == TB_K_PAR || k == TB_K_IMM; }
    double scalar(int32_t o, const tb_op_t &op) const {
        return TB_OPERAND_KIND(o) == TB_K_PAR ? par[TB_OPERAND_INDEX(o)] : op.imm;
    }
};

/* ---- Taylor1 primitives: coefficient k of the result -----------------------
 * Restated from TaylorSeries.jl (external): the `<op>!(res, a, b, k)` methods
 * that `_replacecalls!` emits (src/parse_eqs.jl:963-1135).                    */

/* mul!(c, a, b, k):  c[k] = a[0]*b[k]; for i = 1:k  c[k] += a[i]*b[k-i]      */
inline double p_mul(const double *a, const double *b, int k) {
    double c = a[0] * b[k];
    for (int i = 1; i <= k; i++) c += a[i] * b[k - i];
    return c;
}
/* sqr!(c, a, k) */
inline double p_sqr(const double *a, int k) {
    if (k == 0) return a[0] * a[0];
    int kodd = k % 2;
    int kend = (k - 2 + kodd) / 2;
    double c = a[0] * a[k];
    for (int i = 1; i <= kend; i++) c += a[i] * a[k - i];
    c = 2 * c;
    if (kodd == 0) c += a[k / 2] * a[k / 2];
    return c;
}
/* div!(c, a, b, k), b[0] != 0 (no leading-zero factorisation) */
inline double p_div(const double *c, const double *a, const double *b, int k) {
    if (k == 0) return a[0] / b[0];
    double s = c[0] * b[k];
    for (int i = 1; i <= k - 1; i++) s += c[i] * b[k - i];
    return (a[k] - s) / b[0];
}
/* div!(c, a::Number, b, k):  a is the constant series [a,0,0,...] */
inline double p_div_scalar_num(const double *c, double a, const double *b, int k) {
    if (k == 0) return a / b[0];
    double s = c[0] * b[k];
    for (int i = 1; i <= k - 1; i++) s += c[i] * b[k - i];
    return -s / b[0];
}
/* sqrt!(c, a, k), a[0] > 0 */
inline double p_sqrt(const double *c, const double *a, int k) {
    if (k == 0) return std::sqrt(a[0]);
    int kodd = k % 2;
    int kend = (k - 2 + kodd) / 2;
    double s = 0.0;
    for (int i = 1; i <= kend; i++) {
        if (i == 1) s = c[i] * c[k - i];
        else s += c[i] * c[k - i];
    }
    double aux = a[k] - 2 * s;
    if (kodd == 0) aux -= c[kend + 1] * c[kend + 1];
    return aux / (2 * c[0]);
}
/* pow!(c, a, aux, r, k), general real exponent, a[0] != 0 */
inline double p_pow(const Math &m, const double *c, const double *a, double r, int k) {
    if (k == 0) return m.pow_(a[0], r);
    double s = r * k * c[0] * a[k];
    for (int i = 1; i <= k - 1; i++) {
        double rr = r * (k - i) - i;
        s += rr * c[i] * a[k - i];
    }
    return s / (k * a[0]);
}
/* exp!(c, a, k) */
inline double p_exp(const Math &m, const double *c, const double *a, int k) {
    if (k == 0) return m.exp_(a[0]);
    double s = k * a[k] * c[0];
    for (int i = 1; i <= k - 1; i++) s += (k - i) * a[k - i] * c[i];
    return s / k;
}
/* log!(c, a, k) */
inline double p_log(const Math &m, const double *c, const double *a, int k) {
    if (k == 0) return m.log_(a[0]);
    if (k == 1) return a[1] / a[0];
    double s = (k - 1) * a[1] * c[k - 1];
    for (int i = 2; i <= k - 1; i++) s += (k - i) * a[i] * c[k - i];
    return (a[k] - s / k) / a[0];
}
/* sincos!(s, c, a, k) */
inline void p_sincos(const Math &m, double *s, double *c, const double *a, int k) {
    if (k == 0) { m.sincos_(a[0], &s[0], &c[0]); return; }
    double x = a[1];
    double sk = x * c[k - 1];
    double ck = -x * s[k - 1];
    for (int i = 2; i <= k; i++) {
        x = i * a[i];
        sk += x * c[k - i];
        ck -= x * s[k - i];
    }
    s[k] = sk / k;
    c[k] = ck / k;
}

/* ---- primitives with an auxiliary series (src/parse_eqs.jl:1013-1046): tan!, tanh! (aux = c^2), atan! (aux = 1 + a^2),
 * asin!, acos! (aux = sqrt(1 - a^2)), sinhcosh! (aux = the companion).  TaylorSeries.jl recurrences (external, recalled
 * like the rest of App. B; checked against closed-form Taylor coefficients in tests/test_primitives.py).        */
/* tan!(c, a, c2, k): c[k] = a[k] + (sum_{i<k} (k-i) a[k-i] c2[i]) / k; then sqr!(c2, c, k) */
inline void p_tan(const Math &m, double *c, double *c2, const double *a, int k, bool hyp) {
    if (k == 0) { c[0] = hyp ? m.tanh_(a[0]) : m.tan_(a[0]); c2[0] = c[0] * c[0]; return; }
    double s = k * a[k] * c2[0];
    for (int i = 1; i <= k - 1; i++) s += (k - i) * a[k - i] * c2[i];
    c[k] = hyp ? a[k] - s / k : a[k] + s / k;
    c2[k] = p_sqr(c, k);
}
/* sinhcosh!(s, c, a, k) */
inline void p_sinhcosh(const Math &m, double *s, double *c, const double *a, int k) {
    if (k == 0) { m.sinhcosh_(a[0], &s[0], &c[0]); return; }
    double x = a[1];
    double sk = x * c[k - 1], ck = x * s[k - 1];
    for (int i = 2; i <= k; i++) {
        x = i * a[i];
        sk += x * c[k - i];
        ck += x * s[k - i];
    }
    s[k] = sk / k;
    c[k] = ck / k;
}
/* atan!(c, a, r, k): r = 1 + a^2;  c[k] = (a[k] - (sum_{i=1..k-1} (k-i) r[i] c[k-i]) / k) / r[0] */
inline void p_atan(const Math &m, double *c, double *r, const double *a, int k) {
    if (k == 0) { c[0] = m.atan_(a[0]); r[0] = 1 + a[0] * a[0]; return; }
    double s = (k - 1) * r[1] * c[k - 1];
    for (int i = 2; i <= k - 1; i++) s += (k - i) * r[i] * c[k - i];
    r[k] = p_sqr(a, k);
    c[k] = (a[k] - s / k) / r[0];
}
/* asin!(c, a, r, k) / acos!: r = sqrt(1 - a^2), r[k] from sqrt!(r, 1 - a^2, k) */
inline void p_asin(const Math &m, double *c, double *r, const double *a, int k, bool is_acos) {
    if (k == 0) { c[0] = is_acos ? m.acos_(a[0]) : m.asin_(a[0]); r[0] = std::sqrt(1 - a[0] * a[0]); return; }
    double s = (k - 1) * r[1] * c[k - 1];
    for (int i = 2; i <= k - 1; i++) s += (k - i) * r[i] * c[k - i];
    {   /* sqrt!(r, 1 - a^2, k): coefficient k of 1 - a^2 is -(a^2)[k] */
        const double ak = -p_sqr(a, k);
        const int kodd = k % 2, kend = (k - 2 + kodd) / 2;
        double q = 0.0;
        for (int i = 1; i <= kend; i++) { if (i == 1) q = r[i] * r[k - i]; else q += r[i] * r[k - i]; }
        double aux = ak - 2 * q;
        if (kodd == 0) aux -= r[kend + 1] * r[kend + 1];
        r[k] = aux / (2 * r[0]);
    }
    c[k] = is_acos ? -(a[k] + s / k) / r[0] : (a[k] - s / k) / r[0];
}

int run_op(Work &w, const tb_op_t &op, int k) {
    double *res = w.series(op.res);
    if (!res) return TB_ERR_ARG;
    const bool sa = w.is_scalar(op.a), sb = w.is_scalar(op.b);
    const double *a = sa ? nullptr : w.series(op.a);
    const double *b = sb ? nullptr : w.series(op.b);
    switch (op.op) {
        case TB_OP_IDENTITY: /* identity!(c, a, k) */
            res[k] = sa ? (k == 0 ? w.scalar(op.a, op) : 0.0) : a[k];
            break;
        case TB_OP_NEG: res[k] = -a[k]; break;
        case TB_OP_ZERO: res[k] = 0.0; break;
        case TB_OP_ONE: res[k] = (k == 0) ? 1.0 : 0.0; break;
        case TB_OP_ADD: /* add!: a number only touches k = 0 */
            if (sa && sb) res[k] = k == 0 ? w.scalar(op.a, op) + w.scalar(op.b, op) : 0.0;
            else if (sa) res[k] = k == 0 ? w.scalar(op.a, op) + b[0] : b[k];
            else if (sb) res[k] = k == 0 ? a[0] + w.scalar(op.b, op) : a[k];
            else res[k] = a[k] + b[k];
            break;
        case TB_OP_SUB:
            if (sa && sb) res[k] = k == 0 ? w.scalar(op.a, op) - w.scalar(op.b, op) : 0.0;
            else if (sa) res[k] = k == 0 ? w.scalar(op.a, op) - b[0] : -b[k];
            else if (sb) res[k] = k == 0 ? a[0] - w.scalar(op.b, op) : a[k];
            else res[k] = a[k] - b[k];
            break;
        case TB_OP_MUL:
            if (sa && sb) res[k] = k == 0 ? w.scalar(op.a, op) * w.scalar(op.b, op) : 0.0;
            else if (sa) res[k] = w.scalar(op.a, op) * b[k]; /* mul!(c, a::Number, b, k) */
            else if (sb) res[k] = a[k] * w.scalar(op.b, op);
            else res[k] = p_mul(a, b, k);
            break;
        case TB_OP_DIV:
            if (sa && sb) res[k] = k == 0 ? w.scalar(op.a, op) / w.scalar(op.b, op) : 0.0;
            else if (sb) res[k] = a[k] / w.scalar(op.b, op);
            else if (sa) res[k] = p_div_scalar_num(res, w.scalar(op.a, op), b, k);
            else res[k] = p_div(res, a, b, k);
            break;
        case TB_OP_SQR: res[k] = p_sqr(a, k); break;
        case TB_OP_SQRT: res[k] = p_sqrt(res, a, k); break;
        case TB_OP_POW: { /* pow!(c, a, aux, r, k) dispatch */
            double r = op.imm;
            if (r == 0.0) res[k] = (k == 0) ? 1.0 : 0.0;
            else if (r == 1.0) res[k] = a[k];
            else if (r == 2.0) res[k] = p_sqr(a, k);
            else if (r == 0.5) res[k] = p_sqrt(res, a, k);
            else res[k] = p_pow(w.math, res, a, r, k);
            break;
        }
        case TB_OP_EXP: res[k] = p_exp(w.math, res, a, k); break;
        case TB_OP_LOG: res[k] = p_log(w.math, res, a, k); break;
        case TB_OP_SINCOS: {
            double *cs = w.series(op.aux);
            if (!cs) return TB_ERR_ARG;
            p_sincos(w.math, res, cs, a, k);
            break;
        }
        case TB_OP_TAN: case TB_OP_TANH: case TB_OP_ATAN: case TB_OP_ASIN: case TB_OP_ACOS: case TB_OP_SINHCOSH: {
            double *ax = w.series(op.aux);
            if (!ax || !a) return TB_ERR_ARG;
            if (op.op == TB_OP_TAN || op.op == TB_OP_TANH) p_tan(w.math, res, ax, a, k, op.op == TB_OP_TANH);
            else if (op.op == TB_OP_ATAN) p_atan(w.math, res, ax, a, k);
            else if (op.op == TB_OP_SINHCOSH) p_sinhcosh(w.math, res, ax, a, k);
            else p_asin(w.math, res, ax, a, k, op.op == TB_OP_ACOS);
            break;
        }
        default: return TB_ERR_UNSUPPORTED;
    }
    return 0;
}

/* jetcoeffs!(::Val{f}, t, x, dx, params, rv): src/parse_eqs.jl:259-264 skeleton,
 * recursion 1315-1346; solcoeff! src/TaylorIntegration.jl:33-40 (x[k] = dx[k-1]/k). */
int jetcoeffs(Work &w) {
    const tb_ode_desc_t *d = w.d;
    for (int ord = 0; ord < w.P; ord++) {
        for (int j = 0; j < d->nops; j++) {
            int rc = run_op(w, d->ops[j], ord);
            if (rc) return rc;
        }
        int ordnext = ord + 1;
        for (int i = 0; i < d->dof; i++) w.X(i)[ordnext] = w.DX(i)[ord] / ordnext;
    }
    return 0;
}

/* a tape segment evaluated as full series on completed jets (event function g,
 * Jacobian temporaries): k = 0..kmax                                           */
int run_segment(Work &w, int first, int count, int kmax) {
    for (int k = 0; k <= kmax; k++)
        for (int j = first; j < first + count; j++) {
            int rc = run_op(w, w.d->ops[j], k);
            if (rc) return rc;
        }
    return 0;
}

/* Julia's min / max PROPAGATE NaN (Base.min(x::Float64, y::Float64)), unlike std::min / fmin: a NaN coefficient makes
 * h NaN, then dt, the state and t0 NaN, and the driver loop ends on `sign*t0 < sign*tmax` being false (stepsize.jl:74, 86). */
inline double jl_min(double a, double b) { return (a != a || b != b) ? std::numeric_limits<double>::quiet_NaN() : std::min(a, b); }
inline double jl_max(double a, double b) { return (a != a || b != b) ? std::numeric_limits<double>::quiet_NaN() : std::max(a, b); }

/* _stepsize: src/integrator/stepsize.jl:107-111 */
inline double stepsize_k(const Math &m, double aux1, double epsilon, int k) {
    double aux = epsilon / aux1;
    double kinv = 1.0 / k;
    return m.pow_(aux, kinv);
}
/* stepsize(x::Taylor1, abs, rel): stepsize.jl:58-77 */
double stepsize1(const Math &m, const double *x, int P, double abseps, double releps) {
    double x0norm = std::fabs(x[0]);
    double h = std::numeric_limits<double>::infinity();
    for (int k = P - 1; k <= P; k++) {
        double aux = std::fabs(x[k]);
        if (aux == 0.0) continue;
        double aux1;
        if (abseps >= releps * x0norm) aux1 = stepsize_k(m, aux, abseps, k);
        else aux1 = stepsize_k(m, aux, releps * x0norm, k);
        h = jl_min(h, aux1);
    }
    return h;
}
/* _second_stepsize: stepsize.jl:119-132 */
double second_stepsize(const Math &m, const double *x, int P) {
    bool allzero = true;
    for (int k = 0; k <= P; k++) if (x[k] != 0.0) { allzero = false; break; }
    if (allzero) return std::numeric_limits<double>::infinity();
    double h = 0.0;
    for (int k = 1; k <= P - 2; k++) {
        double aux = std::fabs(x[k]);
        if (aux == 0.0) continue;
        h = jl_max(h, stepsize_k(m, aux, 1.0, k));
    }
    return h;
}
/* stepsize(q::AbstractArray{Taylor1}, abs, rel): stepsize.jl:79-99 */
double stepsize_vec(Work &w, int ncomp, double abseps, double releps) {
    double h = std::numeric_limits<double>::infinity();
    for (int i = 0; i < ncomp; i++) h = jl_min(h, stepsize1(w.math, w.X(i), w.P, abseps, releps));
    if (std::isinf(h)) {
        h = 0.0;
        for (int i = 0; i < ncomp; i++) h = jl_max(h, second_stepsize(w.math, w.X(i), w.P));
    }
    return h;
}
inline double clamp_(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* taylorstep!: src/integrator/taylorstep.jl:59-82 */
int taylorstep(Work &w, const tb_opts_t &o, double *dt) {
    int rc = jetcoeffs(w);
    if (rc) return rc;
    double h = stepsize_vec(w, w.d->dof, o.abstol, o.reltol);
    *dt = clamp_(h, o.minstepsize, o.maxstepsize);
    return 0;
}

/* evaluate(a::Taylor1, dx): Horner, TaylorSeries (external) */
inline double horner(const double *a, int P, double dt) {
    double s = a[P];
    for (int k = P - 1; k >= 0; k--) s = s * dt + a[k];
    return s;
}

/* update_cache!: src/integrator/cache.jl:435-442; init_expansions cache.jl:176-183 */
void update_cache(Work &w, double t0, const double *x0, int ncomp) {
    w.t[0] = t0;
    if (w.P >= 1) w.t[1] = 1.0;
    for (int i = 0; i < ncomp; i++) w.X(i)[0] = x0[i];
}

inline double sgn(double v) { return std::copysign(1.0, v); }

struct StepOut { /* optional per-step storage, column s of trajectory-local arrays */
    double *tv;   /* [maxsteps+1]          */
    double *xv;   /* [dof x (maxsteps+1)]  */
    double *psol; /* [dof x (P+1) x (maxsteps+1)] */
    int64_t stride;
};

/* taylorinteg!(dense, f!, q0, t0, tmax, abstol, cache, params; ...):
 * src/integrator/taylorinteg.jl:329-381                                       */
int integ_plain(Work &w, const double *q0, double t0, double tmax, const tb_opts_t &o, const StepOut *so,
                double *xfinal, double *tfinal, int64_t *nsteps_out, int32_t *status) {
    const int dof = w.d->dof, P = w.P;
    std::vector<double> x0(q0, q0 + dof);
    update_cache(w, t0, x0.data(), dof);
    if (so && so->tv) so->tv[0] = t0;
    if (so && so->xv) for (int i = 0; i < dof; i++) so->xv[(size_t)i * so->stride] = q0[i];
    double sign_tstep = sgn(tmax - t0);
    int64_t nsteps = 1;
    int32_t st = 0;
    while (sign_tstep * t0 < sign_tstep * tmax) {
        double dt;
        int rc = taylorstep(w, o, &dt);
        if (rc) return rc;
        if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
        dt = sign_tstep * jl_min(dt, sign_tstep * (tmax - t0));
        for (int i = 0; i < dof; i++) x0[i] = horner(w.X(i), P, dt);
        if (so && so->psol)
            for (int i = 0; i < dof; i++)
                for (int k = 0; k <= P; k++)
                    so->psol[((size_t)i * (P + 1) + k) * so->stride + (nsteps - 1)] = w.X(i)[k];
        t0 += dt;
        update_cache(w, t0, x0.data(), dof);
        nsteps += 1;
        if (so && so->tv) so->tv[nsteps - 1] = t0;
        if (so && so->xv) for (int i = 0; i < dof; i++) so->xv[(size_t)i * so->stride + (nsteps - 1)] = x0[i];
        if (nsteps > o.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
        bool fin = true;
        for (int i = 0; i < dof; i++) fin = fin && std::isfinite(x0[i]);
        if (!fin) { st |= TB_ST_NONFINITE; break; }
    }
    for (int i = 0; i < dof; i++) xfinal[i] = x0[i];
    *tfinal = t0;
    *nsteps_out = nsteps - 1;
    *status = st;
    return 0;
}

/* taylorinteg!(f!, q0, trange, abstol, cache, params; ...): taylorinteg.jl:582-642 */
int integ_trange(Work &w, const double *q0, const double *trange, int64_t nt, const tb_opts_t &o, double *xv /*[dof x nt]*/,
                 double *xfinal, double *tfinal, int64_t *nsteps_out, int32_t *status) {
    const int dof = w.d->dof, P = w.P;
    const double nan = std::numeric_limits<double>::quiet_NaN();
    for (int64_t j = 0; j < (int64_t)dof * nt; j++) xv[j] = nan; /* cache.jl:313-316 */
    double t0 = trange[0], t1 = trange[1], tmax = trange[nt - 1];
    double sign_tstep = sgn(tmax - t0);
    std::vector<double> x0(q0, q0 + dof), x1(dof);
    update_cache(w, t0, x0.data(), dof);
    for (int i = 0; i < dof; i++) xv[(size_t)i * nt] = q0[i];
    int64_t iter = 2, nsteps = 1; /* 1-based like the reference */
    int32_t st = 0;
    while (sign_tstep * t0 < sign_tstep * tmax) {
        double dt;
        int rc = taylorstep(w, o, &dt);
        if (rc) return rc;
        if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
        dt = sign_tstep * jl_min(dt, sign_tstep * (tmax - t0));
        for (int i = 0; i < dof; i++) x0[i] = horner(w.X(i), P, dt);
        double tnext = t0 + dt;
        while (sign_tstep * t1 < sign_tstep * tnext) {
            for (int i = 0; i < dof; i++) x1[i] = horner(w.X(i), P, t1 - t0);
            for (int i = 0; i < dof; i++) xv[(size_t)i * nt + (iter - 1)] = x1[i];
            iter += 1;
            t1 = trange[iter - 1];
        }
        if (dt == tmax - t0) {
            for (int i = 0; i < dof; i++) xv[(size_t)i * nt + (iter - 1)] = x0[i];
            t0 = tnext;
            break;
        }
        t0 = tnext;
        update_cache(w, t0, x0.data(), dof);
        nsteps += 1;
        if (nsteps > o.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
        bool fin = true;
        for (int i = 0; i < dof; i++) fin = fin && std::isfinite(x0[i]);
        if (!fin) { st |= TB_ST_NONFINITE; break; }
    }
    for (int i = 0; i < dof; i++) xfinal[i] = x0[i];
    *tfinal = t0;
    *nsteps_out = nsteps - 1;
    *status = st;
    return 0;
}

/* Event function g evaluated on the current jets: returns gate flag and fills gser[0..P] */
int eval_event(Work &w, std::vector<double> &gser, bool *flag) {
    const tb_ode_desc_t *d = w.d;
    int rc = run_segment(w, d->nops, d->ev_nops, w.P);
    if (rc) return rc;
    const double *g = w.series(d->ev_g);
    if (!g) return TB_ERR_ARG;
    for (int k = 0; k <= w.P; k++) gser[k] = g[k];
    bool f = true;
    if (d->ev_gate_cmp != TB_GATE_ALWAYS) {
        const double *gs = w.series(d->ev_gate);
        if (!gs) return TB_ERR_ARG;
        double v = gs[0];
        switch (d->ev_gate_cmp) {
            case TB_GATE_GT0: f = v > 0; break;
            case TB_GATE_LT0: f = v < 0; break;
            case TB_GATE_GE0: f = v >= 0; break;
            case TB_GATE_LE0: f = v <= 0; break;
        }
    }
    *flag = f;
    return 0;
}

/* taylorinteg!(dense, f!, g, q0, t0, tmax, abstol, cache, params; ...):
 * src/rootfinding.jl:347-438 with findroot! 176-235, surfacecrossing 124-132,
 * nrconvergencecriterion 147-152.                                             */
int integ_events(Work &w, const double *q0, double t0, double tmax, const tb_opts_t &o, const StepOut *so,
                 double *xfinal, double *tfinal, int64_t *nsteps_out, int32_t *status,
                 double *tevents, double *xevents /*[dof x cap]*/, double *gresids, int32_t *nevents_out, int cap,
                 const double *trange = nullptr, int64_t nt = 0, double *xvtr = nullptr /*[dof x nt]*/) {
    const int dof = w.d->dof, P = w.P, eo = o.eventorder;
    /* trange variant: src/rootfinding.jl:485-584 (in-step outputs 541-546, early exit on the last step 547-550) */
    double t1 = 0.0;
    int64_t iter = 2;
    if (trange) {
        t0 = trange[0]; t1 = trange[1]; tmax = trange[nt - 1];
        const double nan = std::numeric_limits<double>::quiet_NaN();
        for (int64_t j = 0; j < (int64_t)dof * nt; j++) xvtr[j] = nan;
        for (int i = 0; i < dof; i++) xvtr[(size_t)i * nt] = q0[i];
    }
    std::vector<double> x0(q0, q0 + dof), x1(dof);
    update_cache(w, t0, x0.data(), dof);
    if (so && so->tv) so->tv[0] = t0;
    if (so && so->xv) for (int i = 0; i < dof; i++) so->xv[(size_t)i * so->stride] = q0[i];
    double sign_tstep = sgn(tmax - t0);
    /* The reference probes the parsed method on the cache's own x/dx
     * (jetcoeffs.jl:181-182 via cache.jl:266), so the jets are already
     * populated when g is first evaluated (rootfinding.jl:377-378).           */
    int rc = jetcoeffs(w);
    if (rc) return rc;
    std::vector<double> g_old(P + 1), g_now(P + 1), G(P + 1), dG(P + 1);
    bool f_old, f_now;
    rc = eval_event(w, g_old, &f_old);
    if (rc) return rc;
    double dt = 0.0, dt_old = 0.0;
    int64_t nsteps = 1;
    int nevents = 1;
    int32_t st = 0;
    while (sign_tstep * t0 < sign_tstep * tmax) {
        dt_old = dt;
        rc = taylorstep(w, o, &dt);
        if (rc) return rc;
        dt = sign_tstep * jl_min(dt, sign_tstep * (tmax - t0));
        for (int i = 0; i < dof; i++) x0[i] = horner(w.X(i), P, dt);
        if (trange) {
            const double tnext = t0 + dt;
            while (sign_tstep * t1 < sign_tstep * tnext) {
                for (int i = 0; i < dof; i++) x1[i] = horner(w.X(i), P, t1 - t0);
                for (int i = 0; i < dof; i++) xvtr[(size_t)i * nt + (iter - 1)] = x1[i];
                iter += 1;
                t1 = trange[iter - 1];
            }
            if (dt == tmax - t0) {
                for (int i = 0; i < dof; i++) xvtr[(size_t)i * nt + (iter - 1)] = x0[i];
                t0 = tnext;
                break;
            }
        }
        if (so && so->psol)
            for (int i = 0; i < dof; i++)
                for (int k = 0; k <= P; k++)
                    so->psol[((size_t)i * (P + 1) + k) * so->stride + (nsteps - 1)] = w.X(i)[k];
        rc = eval_event(w, g_now, &f_now);
        if (rc) return rc;
        /* findroot! */
        if (f_old && f_now && (g_old[eo] * g_now[eo] < 0.0)) {
            int nriter = 1;
            double slope = (g_now[eo] - g_old[eo]) / dt_old;
            double dt_li = -(g_now[eo] / slope);
            /* derivative(g, eo): repeated d[k] = (k+1) a[k+1]; order stays P, top coefficients zero */
            for (int k = 0; k <= P; k++) G[k] = g_now[k];
            for (int r = 0; r < eo; r++) {
                for (int k = 0; k < P; k++) G[k] = (k + 1) * G[k + 1];
                G[P] = 0.0;
            }
            for (int k = 0; k < P; k++) dG[k] = (k + 1) * G[k + 1];
            dG[P] = 0.0;
            double dt_nr = dt_li;
            double gv = horner(G.data(), P, dt_nr), dgv = horner(dG.data(), P, dt_nr);
            while (std::fabs(gv) > o.nrabstol && nriter <= o.newtoniter) {
                dt_nr = dt_nr - gv / dgv;
                gv = horner(G.data(), P, dt_nr);
                dgv = horner(dG.data(), P, dt_nr);
                nriter += 1;
            }
            if (nriter == o.newtoniter + 1) st |= TB_ST_NR_NOCONV;
            if (nevents <= cap) {
                tevents[nevents - 1] = t0 + dt_nr;
                for (int i = 0; i < dof; i++) xevents[(size_t)i * cap + (nevents - 1)] = horner(w.X(i), P, dt_nr);
                gresids[nevents - 1] = gv;
            } else st |= TB_ST_EVENT_OVERFLOW;
            nevents += 1;
        }
        g_old = g_now;
        f_old = f_now;
        t0 += dt;
        update_cache(w, t0, x0.data(), dof);
        nsteps += 1;
        if (so && so->tv) so->tv[nsteps - 1] = t0;
        if (so && so->xv) for (int i = 0; i < dof; i++) so->xv[(size_t)i * so->stride + (nsteps - 1)] = x0[i];
        if (nsteps > o.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
        bool fin = true;
        for (int i = 0; i < dof; i++) fin = fin && std::isfinite(x0[i]);
        if (!fin) { st |= TB_ST_NONFINITE; break; }
    }
    for (int i = 0; i < dof; i++) xfinal[i] = x0[i];
    *tfinal = t0;
    *nsteps_out = nsteps - 1;
    *status = st;
    *nevents_out = std::min(nevents - 1, cap);
    return 0;
}

/* modifiedGS!(A, Q, R, ...): src/lyapunovspectrum.jl:118-146. Column-major n x n.
 * dot / norm as plain left-to-right sums (LinearAlgebra.generic_norm2 unscaled
 * branch; BLAS ddot order is unpinned).                                        */
void modifiedGS(const double *A, double *Q, double *R, int n) {
    std::vector<double> v(n);
    for (int i = 0; i < n * n; i++) R[i] = 0.0;
    for (int j = 0; j < n; j++) {
        for (int r = 0; r < n; r++) v[r] = A[r + j * n];
        for (int i = 0; i < j; i++) {
            double dot = 0.0;
            for (int r = 0; r < n; r++) {
                if (r == 0) dot = Q[r + i * n] * v[r];
                else dot += Q[r + i * n] * v[r];
            }
            R[i + j * n] = dot;
            for (int r = 0; r < n; r++) v[r] -= dot * Q[r + i * n];
        }
        double ss = 0.0;
        for (int r = 0; r < n; r++) {
            if (r == 0) ss = v[r] * v[r];
            else ss += v[r] * v[r];
        }
        double nrm = std::sqrt(ss);
        R[j + j * n] = nrm;
        for (int r = 0; r < n; r++) Q[r + j * n] = v[r] / nrm;
    }
}

/* lyap_taylorinteg!: src/lyapunovspectrum.jl:340-429; lyap_taylorstep! 226-274;
 * lyap_jetcoeffs! 161-211.  State = [q; vec(Phi)] (column-major Phi).          */
int integ_lyap(Work &w, const double *q0, double t0, double tmax, const tb_opts_t &o, const StepOut *so,
               double *lambda_v /*[dof x stride] or null*/, double *xfinal, double *tfinal, double *lambda,
               int64_t *nsteps_out, int32_t *status, const double *trange = nullptr, int64_t nt = 0,
               double *xvtr = nullptr /*[dof x nt]*/, double *lamtr = nullptr /*[dof x nt]*/,
               const double *state0 = nullptr /* continue from state: Q (column-major), lsum[dof], t00 */,
               double *state_out = nullptr) {
    const tb_ode_desc_t *d = w.d;
    const int dof = d->dof, P = w.P, nx = dof + dof * dof;
    if (!d->has_jac) return TB_ERR_ARG;
    /* trange variant: src/lyapunovspectrum.jl:481-596 (outputs NaN-prefilled, cache.jl:421-422) */
    double t1 = 0.0;
    int64_t iter = 2;
    if (trange) {
        t0 = trange[0]; t1 = trange[1]; tmax = trange[nt - 1];
        const double nan = std::numeric_limits<double>::quiet_NaN();
        for (int64_t j = 0; j < (int64_t)dof * nt; j++) { xvtr[j] = nan; lamtr[j] = nan; }
        for (int i = 0; i < dof; i++) { xvtr[(size_t)i * nt] = q0[i]; lamtr[(size_t)i * nt] = 0.0; }
    }
    std::vector<double> x0(nx, 0.0), jt(dof * dof, 0.0), QH(dof * dof), RH(dof * dof), lsum(dof, 0.0), lam(dof, 0.0);
    for (int i = 0; i < dof; i++) x0[i] = q0[i];
    for (int i = 0; i < dof; i++) jt[i + i * dof] = 1.0;
    if (state0) {
        for (int i = 0; i < dof * dof; i++) jt[i] = state0[i];
        for (int i = 0; i < dof; i++) lsum[i] = state0[dof * dof + i];
    }
    for (int i = 0; i < dof * dof; i++) x0[dof + i] = jt[i];
    /* jac series storage: [dof x dof][P+1] (row i, col j) */
    std::vector<double> jac((size_t)dof * dof * (P + 1), 0.0), vaux(P + 1);
    update_cache(w, t0, x0.data(), nx);
    double t00 = state0 ? state0[dof * dof + dof] : t0;
    if (so && so->tv) so->tv[0] = t0;
    if (so && so->xv) for (int i = 0; i < dof; i++) so->xv[(size_t)i * so->stride] = q0[i];
    if (lambda_v) for (int i = 0; i < dof; i++) lambda_v[(size_t)i * so->stride] = 0.0;
    double sign_tstep = sgn(tmax - t0);
    int64_t nsteps = 1;
    int32_t st = 0;
    while (sign_tstep * t0 < sign_tstep * tmax) {
        /* lyap_taylorstep!: trajectory jets */
        int rc = jetcoeffs(w);
        if (rc) return rc;
        /* stabilitymatrix!: order-1 TaylorN pass done at tape-build time; evaluate its series */
        rc = run_segment(w, d->nops + d->ev_nops, d->jac_nops, P - 1);
        if (rc) return rc;
        for (int i = 0; i < dof; i++)
            for (int j = 0; j < dof; j++) {
                double *J = &jac[((size_t)i * dof + j) * (P + 1)];
                const tb_jac_entry_t &e = d->jac[i * dof + j];
                for (int k = 0; k <= P; k++) J[k] = 0.0;
                if (e.kind == TB_JAC_CONST) J[0] = e.c;
                else if (e.kind == TB_JAC_PAR) J[0] = e.c * w.par[TB_OPERAND_INDEX(e.operand)];
                else if (e.kind == TB_JAC_SERIES) {
                    const double *s = w.series(e.operand);
                    if (!s) return TB_ERR_ARG;
                    for (int k = 0; k < P; k++) J[k] = s[k];
                }
            }
        /* lyap_jetcoeffs!: dx = jac * Phi, CartesianIndex order (i1 fastest, i2, i3 slowest) */
        for (int idx = 0; idx < dof * dof; idx++)
            for (int k = 0; k <= P; k++) w.DX(dof + idx)[k] = 0.0;
        for (int i3 = 0; i3 < dof; i3++)
            for (int i2 = 0; i2 < dof; i2++)
                for (int i1 = 0; i1 < dof; i1++) {
                    const double *J = &jac[((size_t)i1 * dof + i3) * (P + 1)];
                    double v = J[0] * w.X(dof + i3 + i2 * dof)[0];
                    double *D = w.DX(dof + i1 + i2 * dof);
                    D[0] = D[0] + v;
                }
        for (int idx = 0; idx < dof * dof; idx++) w.X(dof + idx)[1] = w.DX(dof + idx)[0];
        for (int ord = 1; ord < P; ord++) {
            int ordnext = ord + 1;
            for (int i3 = 0; i3 < dof; i3++)
                for (int i2 = 0; i2 < dof; i2++)
                    for (int i1 = 0; i1 < dof; i1++) {
                        const double *J = &jac[((size_t)i1 * dof + i3) * (P + 1)];
                        double v = p_mul(J, w.X(dof + i3 + i2 * dof), ord);
                        double *D = w.DX(dof + i1 + i2 * dof);
                        D[ord] = D[ord] + v;
                    }
            for (int idx = 0; idx < dof * dof; idx++) w.X(dof + idx)[ordnext] = w.DX(dof + idx)[ord] / ordnext;
        }
        double dt = stepsize_vec(w, dof, o.abstol, o.reltol);
        dt = clamp_(dt, o.minstepsize, o.maxstepsize);
        dt = sign_tstep * jl_min(dt, sign_tstep * (tmax - t0));
        for (int i = 0; i < nx; i++) x0[i] = horner(w.X(i), P, dt);
        for (int i = 0; i < dof * dof; i++) jt[i] = x0[dof + i];
        modifiedGS(jt.data(), QH.data(), RH.data(), dof);
        if (trange) {
            /* in-step outputs (lyapunovspectrum.jl:545-558): the state at t1, the exponents from the fundamental
             * matrix evaluated at the END of the step (the reference evaluates x[dof+ind] at dt, not at t1 - t0), and
             * the loop test is the reference's literal `t1 < tnext` (no sign factor)                               */
            const double tnext = t0 + dt;
            while (t1 < tnext) {
                for (int i = 0; i < dof; i++) xvtr[(size_t)i * nt + (iter - 1)] = horner(w.X(i), P, t1 - t0);
                const double tsp = t1 - t00;
                for (int i = 0; i < dof; i++) lamtr[(size_t)i * nt + (iter - 1)] = (lsum[i] + w.math.log_(RH[i + i * dof])) / tsp;
                iter += 1;
                t1 = trange[iter - 1];
            }
            if (dt == tmax - t0) {
                const double tsp = tmax - t00;
                for (int i = 0; i < dof; i++) {
                    xvtr[(size_t)i * nt + (iter - 1)] = x0[i];
                    lam[i] = (lsum[i] + w.math.log_(RH[i + i * dof])) / tsp;
                    lamtr[(size_t)i * nt + (iter - 1)] = lam[i];
                }
                t0 = tnext;
                break;
            }
        }
        t0 += dt;
        double tspan = t0 - t00;
        nsteps += 1;
        if (so && so->tv) so->tv[nsteps - 1] = t0;
        for (int i = 0; i < dof; i++) {
            if (so && so->xv) so->xv[(size_t)i * so->stride + (nsteps - 1)] = x0[i];
            lsum[i] += w.math.log_(RH[i + i * dof]);
            lam[i] = lsum[i] / tspan;
            if (lambda_v) lambda_v[(size_t)i * so->stride + (nsteps - 1)] = lam[i];
        }
        for (int i = 0; i < dof * dof; i++) x0[dof + i] = QH[i];
        update_cache(w, t0, x0.data(), nx);
        if (nsteps > o.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
        bool fin = true;
        for (int i = 0; i < dof; i++) fin = fin && std::isfinite(x0[i]);
        if (!fin) { st |= TB_ST_NONFINITE; break; }
    }
    for (int i = 0; i < dof; i++) { xfinal[i] = x0[i]; lambda[i] = lam[i]; }
    if (state_out) {
        for (int i = 0; i < dof * dof; i++) state_out[i] = x0[dof + i];
        for (int i = 0; i < dof; i++) state_out[dof * dof + i] = lsum[i];
        state_out[dof * dof + dof] = t00;
    }
    *tfinal = t0;
    *nsteps_out = nsteps - 1;
    *status = st;
    return 0;
}

void load_params(Work &w, const double *params, int64_t N, int64_t i, int per_traj) {
    for (int j = 0; j < w.d->npar; j++) w.par[j] = per_traj ? params[(size_t)j * N + i] : params[j];
}

template <class F>
void parallel_for(int64_t N, int nthreads, F f) {
    if (nthreads <= 1 || N < 2) { f(0, N, 0); return; }
    std::vector<std::thread> th;
    int64_t chunk = (N + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        int64_t lo = t * chunk, hi = std::min<int64_t>(N, lo + chunk);
        if (lo >= hi) break;
        th.emplace_back([=] { f(lo, hi, t); });
    }
    for (auto &x : th) x.join();
}

} // namespace

extern "C" {

/* modifiedGS! on one column-major n x n matrix (tests: the properties test/lyapunov.jl:88-110 asks of the QR step) */
void tbo_modified_gs(const double *A, double *Q, double *R, int n) { modifiedGS(A, Q, R, n); }

/* accuracy probes for tests/test_math.py */
double tbo_math_pow(double x, double y) { return tb_pow(x, y); }
void tbo_math_pow_pair(double x1, double y1, double x2, double y2, double *r) { tb_pair p = tb_pow_pair(x1, y1, x2, y2); r[0] = p.a; r[1] = p.b; }
double tbo_math_exp(double x) { return tb_exp(x); }
double tbo_math_log(double x) { return tb_log(x); }
void tbo_math_sincos(double x, double *s, double *c) { tb_sincos(x, s, c); }
double tbo_math_tan(double x) { return tb_tan(x); }
double tbo_math_tanh(double x) { return tb_tanh(x); }
double tbo_math_atan(double x) { return tb_atan(x); }
double tbo_math_asin(double x) { return tb_asin(x); }
double tbo_math_acos(double x) { return tb_acos(x); }
void tbo_math_sinhcosh(double x, double *s, double *c) { tb_sinhcosh(x, s, c); }

/* First step: jets x[dof x (P+1)] (row per component), dx likewise, and dt.
 * kind = TB_KIND_LYAP is not supported here (use tbo_integrate).              */
int tbo_jet(const tb_ode_desc_t *d, int order, int math_mode, const double *q0, const double *params, double t0,
            const tb_opts_t *o, double *xjet, double *dxjet, double *dt) {
    Work w(d, order, math_mode);
    for (int j = 0; j < d->npar; j++) w.par[j] = params[j];
    update_cache(w, t0, q0, d->dof);
    int rc = taylorstep(w, *o, dt);
    if (rc) return rc;
    for (int i = 0; i < d->dof; i++)
        for (int k = 0; k <= order; k++) {
            if (xjet) xjet[(size_t)i * (order + 1) + k] = w.X(i)[k];
            if (dxjet) dxjet[(size_t)i * (order + 1) + k] = w.DX(i)[k];
        }
    return 0;
}

/* Series of any operand after one jetcoeffs! call (tape debugging / AD tests):
 * evaluates the RHS tape and optionally the Jacobian segment, returns series
 * of `operand` in out[0..order].                                              */
int tbo_series(const tb_ode_desc_t *d, int order, int math_mode, const double *q0, const double *params, double t0,
               int32_t operand, int with_jac, double *out) {
    Work w(d, order, math_mode);
    for (int j = 0; j < d->npar; j++) w.par[j] = params[j];
    update_cache(w, t0, q0, d->dof);
    int rc = jetcoeffs(w);
    if (rc) return rc;
    if (with_jac) { rc = run_segment(w, d->nops + d->ev_nops, d->jac_nops, order - 1); if (rc) return rc; }
    if (d->ev_nops) { rc = run_segment(w, d->nops, d->ev_nops, order); if (rc) return rc; }
    const double *s = w.series(operand);
    if (!s) return TB_ERR_ARG;
    for (int k = 0; k <= order; k++) out[k] = s[k];
    return 0;
}

/* Ensemble driver with the same SoA layout and outputs as tb_integrate
 * (include/taylorb200.h); `kind` selects the reference driver, `nthreads`
 * OS threads with static chunking stand in for a Threads.@threads loop over
 * taylorinteg calls (docs/src/root_finding.md:121-132 is the serial idiom).   */
int tbo_integrate(const tb_ode_desc_t *d, int order, int kind, int math_mode, int nthreads, int64_t N,
                  const double *q0, const double *params, double t0, double tmax, const double *trange, int64_t nt,
                  const tb_opts_t *opts, tb_out_t *out) {
    if (!d || !opts || !out || N < 0) return TB_ERR_ARG;
    const int dof = d->dof;
    const tb_opts_t o = *opts;
    const int64_t stride = o.steps_stride ? o.steps_stride : o.maxsteps + 1;
    const int cap = o.event_capacity > 0 ? o.event_capacity : (int)(o.maxsteps + 1);
    std::vector<int> rcs(std::max(1, nthreads), 0);
    parallel_for(N, nthreads, [&](int64_t lo, int64_t hi, int tid) {
        Work w(d, order, math_mode, kind == TB_KIND_LYAP ? dof * dof : 0);
        std::vector<double> q(dof), xf(dof), lam(dof);
        std::vector<double> tvl, xvl, psl, lvl, xtr, tev, xev, gre;
        const bool want_steps = !trange && (out->tv || out->xv || out->psol || out->lambda_v);
        if (want_steps) {
            tvl.assign(stride, 0.0);
            xvl.assign((size_t)dof * stride, 0.0);
            if (out->psol) psl.assign((size_t)dof * (order + 1) * stride, 0.0);
            if (out->lambda_v) lvl.assign((size_t)dof * stride, 0.0);
        }
        std::vector<double> ltr;
        if (trange) { xtr.assign((size_t)dof * nt, 0.0); ltr.assign((size_t)dof * nt, 0.0); }
        if (kind == TB_KIND_EVENTS) { tev.assign(cap, 0.0); xev.assign((size_t)dof * cap, 0.0); gre.assign(cap, 0.0); }
        for (int64_t i = lo; i < hi; i++) {
            for (int c = 0; c < dof; c++) q[c] = q0[(size_t)c * N + i];
            load_params(w, params, N, i, o.params_per_traj);
            StepOut so{want_steps ? tvl.data() : nullptr, want_steps ? xvl.data() : nullptr,
                       out->psol ? psl.data() : nullptr, stride};
            double tf = t0;
            int64_t ns = 0;
            int32_t st = 0, nev = 0;
            int rc = 0;
            if (trange && kind == TB_KIND_LYAP)
                rc = integ_lyap(w, q.data(), t0, tmax, o, nullptr, nullptr, xf.data(), &tf, lam.data(), &ns, &st, trange, nt,
                                xtr.data(), ltr.data());
            else if (trange && kind == TB_KIND_EVENTS)
                rc = integ_events(w, q.data(), t0, tmax, o, nullptr, xf.data(), &tf, &ns, &st, tev.data(), xev.data(),
                                  gre.data(), &nev, cap, trange, nt, xtr.data());
            else if (trange) rc = integ_trange(w, q.data(), trange, nt, o, xtr.data(), xf.data(), &tf, &ns, &st);
            else if (kind == TB_KIND_EVENTS)
                rc = integ_events(w, q.data(), t0, tmax, o, want_steps ? &so : nullptr, xf.data(), &tf, &ns, &st,
                                  tev.data(), xev.data(), gre.data(), &nev, cap);
            else if (kind == TB_KIND_LYAP) {
                const int rows = dof * dof + dof + 1;
                std::vector<double> s0, s1(rows);
                if (out->lyap_state0) { s0.resize(rows); for (int r = 0; r < rows; r++) s0[r] = out->lyap_state0[(size_t)r * N + i]; }
                rc = integ_lyap(w, q.data(), t0, tmax, o, want_steps ? &so : nullptr,
                                out->lambda_v ? lvl.data() : nullptr, xf.data(), &tf, lam.data(), &ns, &st, nullptr, 0, nullptr,
                                nullptr, out->lyap_state0 ? s0.data() : nullptr, s1.data());
                if (out->lyap_state) for (int r = 0; r < rows; r++) out->lyap_state[(size_t)r * N + i] = s1[r];
            }

            else rc = integ_plain(w, q.data(), t0, tmax, o, want_steps ? &so : nullptr, xf.data(), &tf, &ns, &st);
            if (rc) { rcs[tid] = rc; return; }
            if (out->xfinal) for (int c = 0; c < dof; c++) out->xfinal[(size_t)c * N + i] = xf[c];
            if (out->tfinal) out->tfinal[i] = tf;
            if (out->nsteps) out->nsteps[i] = ns;
            if (out->status) out->status[i] = st;
            if (out->lambda && kind == TB_KIND_LYAP) for (int c = 0; c < dof; c++) out->lambda[(size_t)c * N + i] = lam[c];
            if (trange && out->xv_trange)
                for (int c = 0; c < dof; c++)
                    for (int64_t j = 0; j < nt; j++) out->xv_trange[((size_t)c * nt + j) * N + i] = xtr[(size_t)c * nt + j];
            if (trange && kind == TB_KIND_LYAP && out->lambda_v)
                for (int c = 0; c < dof; c++)
                    for (int64_t j = 0; j < nt; j++) out->lambda_v[((size_t)c * nt + j) * N + i] = ltr[(size_t)c * nt + j];
            if (want_steps) {
                for (int64_t s = 0; s <= ns && s < stride; s++) {
                    if (out->tv) out->tv[(size_t)s * N + i] = tvl[s];
                    for (int c = 0; c < dof; c++) {
                        if (out->xv) out->xv[((size_t)c * stride + s) * N + i] = xvl[(size_t)c * stride + s];
                        if (out->lambda_v) out->lambda_v[((size_t)c * stride + s) * N + i] = lvl[(size_t)c * stride + s];
                    }
                }
                if (out->psol)
                    for (int c = 0; c < dof; c++)
                        for (int k = 0; k <= order; k++)
                            for (int64_t s = 0; s < ns && s < stride; s++)
                                out->psol[(((size_t)c * (order + 1) + k) * stride + s) * N + i] =
                                    psl[((size_t)c * (order + 1) + k) * stride + s];
            }
            if (kind == TB_KIND_EVENTS) {
                if (out->nevents) out->nevents[i] = nev;
                for (int e = 0; e < nev; e++) {
                    if (out->tevents) out->tevents[(size_t)e * N + i] = tev[e];
                    if (out->gresids) out->gresids[(size_t)e * N + i] = gre[e];
                    if (out->xevents) for (int c = 0; c < dof; c++) out->xevents[((size_t)c * cap + e) * N + i] = xev[(size_t)c * cap + e];
                }
            }
        }
    });
    for (int rc : rcs) if (rc) return rc;
    return 0;
}

int tbo_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

} /* extern "C" */
