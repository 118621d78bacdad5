/*
 * tb_generic.cpp -- the reference's GENERIC jetcoeffs! path, restated on the CPU.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/tb_oracle.cpp).  Nothing in the product links or calls this file.
 *
 * src/integrator/jetcoeffs.jl:66-92: for ord = 0..order-1 the user's RHS is RE-EVALUATED on copies of the jets
 * truncated to order `ord` (`xaux[j] = Taylor1(x[j].coeffs[1:ord+1])`) with ordinary full-series Taylor1 arithmetic
 * (`*`, `/`, `^`, `sin`, ... each computing ALL coefficients 0..ord), then `solcoeff!` takes coefficient `ord` of dx.
 * That is O(order^3) and shares no code with the tape path (oracle/tb_oracle.cpp: one `<op>!(res, a, b, ord)` call per
 * tape entry and order, from the tape built by taylorintegration.jl_b200/tape.py).  The reference's strongest
 * invariant is that both paths agree BITWISE (test/taylorize.jl:1267-1273, 1354-1364, 1817-1819,
 * `iszero(norm(xv - xvp, Inf))`); tests/test_generic.py asserts the same between this file and the tape oracle, which
 * gives the tape front end (shared by product and oracle) a check that does not pass through it: the right-hand
 * sides below are written by hand from the Julia sources they cite.
 *
 * The Taylor1 value type below restates TaylorSeries.jl 0.22 full-series operations (external dependency, not under
 * /root/reference): each builds its result coefficient by coefficient with the same recurrence as the in-place
 * `<op>!(c, a, b, k)` kernels, left-to-right sums, separate multiply and add.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../include/taylorb200.h"
#include "../taylorintegration.jl_b200/csrc/tb_math.h"

namespace {

struct T1 { /* Taylor1{Float64} */
    std::vector<double> c;
    T1() {}
    explicit T1(int order) : c(order + 1, 0.0) {}
    T1(const double *p, int n) : c(p, p + n) {}
    int order() const { return (int)c.size() - 1; }
};

/* fixorder: binary operations pad the shorter operand with zeros (TaylorSeries `fixorder`) */
inline void fix(const T1 &a, const T1 &b, T1 &A, T1 &B) {
    A = a; B = b;
    const size_t n = std::max(a.c.size(), b.c.size());
    A.c.resize(n, 0.0); B.c.resize(n, 0.0);
}
inline T1 operator+(const T1 &a, const T1 &b) { T1 A, B; fix(a, b, A, B); for (size_t k = 0; k < A.c.size(); k++) A.c[k] = A.c[k] + B.c[k]; return A; }
inline T1 operator-(const T1 &a, const T1 &b) { T1 A, B; fix(a, b, A, B); for (size_t k = 0; k < A.c.size(); k++) A.c[k] = A.c[k] - B.c[k]; return A; }
inline T1 operator-(const T1 &a) { T1 r = a; for (double &v : r.c) v = -v; return r; }
inline T1 operator+(const T1 &a, double s) { T1 r = a; r.c[0] = a.c[0] + s; return r; }
inline T1 operator-(double s, const T1 &a) { T1 r = -a; r.c[0] = s - a.c[0]; return r; }
inline T1 operator-(const T1 &a, double s) { T1 r = a; r.c[0] = a.c[0] - s; return r; }
inline T1 operator*(double s, const T1 &a) { T1 r = a; for (double &v : r.c) v = s * v; return r; }
inline T1 operator*(const T1 &a, double s) { T1 r = a; for (double &v : r.c) v = v * s; return r; }
inline T1 operator/(const T1 &a, double s) { T1 r = a; for (double &v : r.c) v = v / s; return r; }
/* a * b: c[k] = a[0]*b[k] + a[1]*b[k-1] + ... */
inline T1 operator*(const T1 &a, const T1 &b) {
    T1 A, B; fix(a, b, A, B);
    T1 r(A.order());
    for (int k = 0; k <= r.order(); k++) {
        double s = A.c[0] * B.c[k];
        for (int i = 1; i <= k; i++) s += A.c[i] * B.c[k - i];
        r.c[k] = s;
    }
    return r;
}
/* a / b, b[0] != 0 */
inline T1 operator/(const T1 &a, const T1 &b) {
    T1 A, B; fix(a, b, A, B);
    T1 r(A.order());
    r.c[0] = A.c[0] / B.c[0];
    for (int k = 1; k <= r.order(); k++) {
        double s = r.c[0] * B.c[k];
        for (int i = 1; i <= k - 1; i++) s += r.c[i] * B.c[k - i];
        r.c[k] = (A.c[k] - s) / B.c[0];
    }
    return r;
}
inline T1 square(const T1 &a) {
    T1 r(a.order());
    r.c[0] = a.c[0] * a.c[0];
    for (int k = 1; k <= r.order(); k++) {
        const int kodd = k % 2, kend = (k - 2 + kodd) / 2;
        double s = a.c[0] * a.c[k];
        for (int i = 1; i <= kend; i++) s += a.c[i] * a.c[k - i];
        s = 2 * s;
        if (kodd == 0) s += a.c[k / 2] * a.c[k / 2];
        r.c[k] = s;
    }
    return r;
}
inline T1 sqrt_(const T1 &a) {
    T1 r(a.order());
    r.c[0] = std::sqrt(a.c[0]);
    for (int k = 1; k <= r.order(); k++) {
        const int kodd = k % 2, kend = (k - 2 + kodd) / 2;
        double s = 0.0;
        for (int i = 1; i <= kend; i++) { if (i == 1) s = r.c[i] * r.c[k - i]; else s += r.c[i] * r.c[k - i]; }
        double aux = a.c[k] - 2 * s;
        if (kodd == 0) aux -= r.c[kend + 1] * r.c[kend + 1];
        r.c[k] = aux / (2 * r.c[0]);
    }
    return r;
}
/* a ^ r for a real exponent: 0 -> one, 1 -> copy, 2 -> square, 0.5 -> sqrt, else the general recurrence */
inline T1 pow_(const T1 &a, double r) {
    if (r == 0.0) { T1 o(a.order()); o.c[0] = 1.0; return o; }
    if (r == 1.0) return a;
    if (r == 2.0) return square(a);
    if (r == 0.5) return sqrt_(a);
    T1 c(a.order());
    c.c[0] = tb_pow(a.c[0], r);
    for (int k = 1; k <= c.order(); k++) {
        double s = r * k * c.c[0] * a.c[k];
        for (int i = 1; i <= k - 1; i++) { const double rr = r * (k - i) - i; s += rr * c.c[i] * a.c[k - i]; }
        c.c[k] = s / (k * a.c[0]);
    }
    return c;
}
inline void sincos_(const T1 &a, T1 &s, T1 &c) {
    s = T1(a.order()); c = T1(a.order());
    tb_sincos(a.c[0], &s.c[0], &c.c[0]);
    for (int k = 1; k <= a.order(); k++) {
        double x = a.c[1];
        double sk = x * c.c[k - 1], ck = -x * s.c[k - 1];
        for (int i = 2; i <= k; i++) { x = i * a.c[i]; sk += x * c.c[k - i]; ck -= x * s.c[k - i]; }
        s.c[k] = sk / k; c.c[k] = ck / k;
    }
}
inline T1 sin_(const T1 &a) { T1 s, c; sincos_(a, s, c); return s; }
inline T1 cos_(const T1 &a) { T1 s, c; sincos_(a, s, c); return c; }
inline T1 exp_(const T1 &a) {
    T1 c(a.order());
    c.c[0] = tb_exp(a.c[0]);
    for (int k = 1; k <= c.order(); k++) {
        double s = k * a.c[k] * c.c[0];
        for (int i = 1; i <= k - 1; i++) s += (k - i) * a.c[k - i] * c.c[i];
        c.c[k] = s / k;
    }
    return c;
}

/* ---- right-hand sides, written from the Julia sources ---------------------------------------------------- */
typedef void (*Rhs)(std::vector<T1> &dx, const std::vector<T1> &x, const double *p, const T1 &t);

/* benchmark/kepler_benchmarks.jl:9-19 (kepler1!) */
void kepler1(std::vector<T1> &dq, const std::vector<T1> &q, const double *p, const T1 &) {
    const double mu = p[0];
    T1 r_p3d2 = pow_(pow_(q[0], 2.0) + pow_(q[1], 2.0), 1.5);
    dq[0] = q[2];
    dq[1] = q[3];
    dq[2] = (mu * q[0]) / r_p3d2;
    dq[3] = (mu * q[1]) / r_p3d2;
}
/* test/rootfinding.jl:20-24 (pendulum!) */
void pendulum(std::vector<T1> &dx, const std::vector<T1> &x, const double *, const T1 &) {
    dx[0] = x[1];
    dx[1] = -sin_(x[0]);
}
/* benchmark/Lyap_benchmark.jl:3-9 (lorenz!) */
void lorenz(std::vector<T1> &dx, const std::vector<T1> &x, const double *p, const T1 &) {
    const double sigma = p[0], rho = p[1], beta = p[2];
    dx[0] = sigma * (x[1] - x[0]);
    dx[1] = x[0] * (rho - x[2]) - x[1];
    dx[2] = x[0] * x[1] - beta * x[2];
}
/* test/one_ode.jl:13: xdot = x^2 (vector form, test/many_ode.jl:19-24) */
void xsquared(std::vector<T1> &dx, const std::vector<T1> &x, const double *, const T1 &) { dx[0] = pow_(x[0], 2.0); }
/* test/many_ode.jl:239-245: xdot = cos(t) (non-autonomous: the time series is an operand) */
void cos_t(std::vector<T1> &dx, const std::vector<T1> &, const double *, const T1 &t) { dx[0] = cos_(t); }
/* test/taylorize.jl:878-890: harmonic oscillator, dx2 = -(omega^2 * x1), omega^2 a plain number */
void harmonic(std::vector<T1> &dx, const std::vector<T1> &x, const double *p, const T1 &) {
    dx[0] = x[1];
    dx[1] = -((p[0] * p[0]) * x[0]);
}

/* benchmark/manyspin_benchmarks.jl:10-41 (spin_odes!) with nn = 5; params = [B[1..5], J[a, i] at 5 + a*5 + i] */
void spin5(std::vector<T1> &dq, const std::vector<T1> &q, const double *p, const T1 &) {
    const int nn = 5;
    const double *B = p, *J = p + nn;
    std::vector<T1> x(nn), y(nn), z(nn), sum_s(nn), aux(nn);
    for (int i = 0; i < nn; i++) { x[i] = q[3 * i]; y[i] = q[3 * i + 1]; z[i] = q[3 * i + 2]; }
    T1 zz(x[0].order()); /* zero(x[1]) */
    for (int i = 0; i < nn; i++) {
        sum_s[i] = zz;
        for (int a = 0; a < nn; a++) {
            aux[i] = sum_s[i];
            T1 xJ = J[a * nn + i] * x[a];
            sum_s[i] = aux[i] + xJ;
        }
        dq[3 * i] = (-B[i]) * y[i];
        dq[3 * i + 1] = B[i] * x[i] + z[i] * sum_s[i];
        dq[3 * i + 2] = (-y[i]) * sum_s[i];
    }
}

/* test/taylorize.jl:1488-1500: the Kepler problem "using sqrt", r = sqrt(q1^2 + q2^2), r^3, mu = -1 a local */
void kepler_sqrt(std::vector<T1> &dq, const std::vector<T1> &q, const double *, const T1 &) {
    const double mu = -1.0;
    T1 r = sqrt_(pow_(q[0], 2.0) + pow_(q[1], 2.0));
    T1 r_p3d2 = pow_(r, 3.0);
    dq[0] = q[2];
    dq[1] = q[3];
    dq[2] = (mu * q[0]) / r_p3d2;
    dq[3] = (mu * q[1]) / r_p3d2;
}
/* test/taylorize.jl:999-1005 (multpendula1!): 3 uncoupled pendula written as a loop */
void multpendula3(std::vector<T1> &dx, const std::vector<T1> &x, const double *, const T1 &) {
    const int nn = 3;
    for (int i = 0; i < nn; i++) {
        dx[i] = x[nn + i];
        dx[i + nn] = -sin_(x[i]);
    }
}

struct Problem { const char *name; int dof; Rhs f; };
const Problem PROBLEMS[] = {{"kepler1!", 4, kepler1}, {"pendulum!", 2, pendulum}, {"lorenz!", 3, lorenz},
                            {"xsquared", 1, xsquared}, {"cos_t", 1, cos_t}, {"harmonic", 2, harmonic},
                            {"spin_odes!5", 15, spin5}, {"kepler_sqrt!", 4, kepler_sqrt}, {"multpendula!", 6, multpendula3}};
const int NPROBLEMS = (int)(sizeof(PROBLEMS) / sizeof(PROBLEMS[0]));

/* jetcoeffs!(eqsdiff!, t, x, dx, xaux, params): src/integrator/jetcoeffs.jl:66-92 */
void jetcoeffs_generic(const Problem &pb, int P, const T1 &t, std::vector<T1> &x, const double *par) {
    std::vector<T1> dx(pb.dof), xaux(pb.dof);
    for (int ord = 0; ord < P; ord++) {
        const int ordnext = ord + 1;
        for (int j = 0; j < pb.dof; j++) xaux[j] = T1(x[j].c.data(), ordnext);
        pb.f(dx, xaux, par, t);
        for (int j = 0; j < pb.dof; j++) x[j].c[ordnext] = dx[j].c[ord] / ordnext; /* solcoeff!: TaylorIntegration.jl:33-40 */
    }
}

inline double jl_min(double a, double b) { return (a != a || b != b) ? std::numeric_limits<double>::quiet_NaN() : std::min(a, b); }
inline double jl_max(double a, double b) { return (a != a || b != b) ? std::numeric_limits<double>::quiet_NaN() : std::max(a, b); }

/* stepsize(q, abs, rel) + clamp: src/integrator/stepsize.jl:58-132, taylorstep.jl:79 */
double stepsize(const std::vector<T1> &x, int P, const tb_opts_t &o) {
    const double inf = std::numeric_limits<double>::infinity();
    double h = inf;
    for (const T1 &xi : x) {
        const double x0norm = std::fabs(xi.c[0]);
        double hi = inf;
        for (int k = P - 1; k <= P; k++) {
            const double aux = std::fabs(xi.c[k]);
            if (aux == 0.0) continue;
            const double eps = (o.abstol >= o.reltol * x0norm) ? o.abstol : o.reltol * x0norm;
            hi = jl_min(hi, tb_pow(eps / aux, 1.0 / k));
        }
        h = jl_min(h, hi);
    }
    if (std::isinf(h)) {
        h = 0.0;
        for (const T1 &xi : x) {
            bool allzero = true;
            for (double v : xi.c) if (v != 0.0) { allzero = false; break; }
            double hc = inf;
            if (!allzero) {
                hc = 0.0;
                for (int k = 1; k <= P - 2; k++) {
                    const double aux = std::fabs(xi.c[k]);
                    if (aux == 0.0) continue;
                    hc = jl_max(hc, tb_pow(1.0 / aux, 1.0 / k));
                }
            }
            h = jl_max(h, hc);
        }
    }
    return h < o.minstepsize ? o.minstepsize : (h > o.maxstepsize ? o.maxstepsize : h);
}
inline double horner(const T1 &a, double dt) {
    double s = a.c.back();
    for (int k = a.order() - 1; k >= 0; k--) s = s * dt + a.c[k];
    return s;
}

} // namespace

extern "C" {

int tbg_problem_count(void) { return NPROBLEMS; }
const char *tbg_problem_name(int id) { return (id >= 0 && id < NPROBLEMS) ? PROBLEMS[id].name : nullptr; }
int tbg_problem_dof(int id) { return (id >= 0 && id < NPROBLEMS) ? PROBLEMS[id].dof : -1; }

/* one generic jetcoeffs! call: xjet [dof x (order+1)] */
int tbg_jet(int id, int order, const double *q0, const double *params, double t0, double *xjet) {
    if (id < 0 || id >= NPROBLEMS || order < 1) return TB_ERR_ARG;
    const Problem &pb = PROBLEMS[id];
    T1 t(order);
    t.c[0] = t0; t.c[1] = 1.0;
    std::vector<T1> x(pb.dof, T1(order));
    for (int j = 0; j < pb.dof; j++) x[j].c[0] = q0[j];
    jetcoeffs_generic(pb, order, t, x, params);
    for (int j = 0; j < pb.dof; j++) memcpy(xjet + (size_t)j * (order + 1), x[j].c.data(), (size_t)(order + 1) * 8);
    return TB_OK;
}

/* taylorinteg!(..., parse_eqs = false): src/integrator/taylorinteg.jl:329-381 with the generic jetcoeffs!.
 * One trajectory; tv [maxsteps+1], xv [dof x (maxsteps+1)] (either may be NULL), returns nsteps via *nsteps_out. */
int tbg_integrate(int id, int order, const double *q0, const double *params, double t0, double tmax, const tb_opts_t *o,
                  double *tv, double *xv, double *xfinal, double *tfinal, int64_t *nsteps_out, int32_t *status) {
    if (id < 0 || id >= NPROBLEMS || order < 1 || !o) return TB_ERR_ARG;
    const Problem &pb = PROBLEMS[id];
    const int dof = pb.dof;
    const int64_t stride = o->maxsteps + 1;
    T1 t(order);
    t.c[0] = t0; t.c[1] = 1.0;
    std::vector<T1> x(dof, T1(order));
    std::vector<double> x0(q0, q0 + dof);
    for (int j = 0; j < dof; j++) x[j].c[0] = x0[j];
    if (tv) tv[0] = t0;
    if (xv) for (int j = 0; j < dof; j++) xv[(size_t)j * stride] = x0[j];
    const double sign = std::copysign(1.0, tmax - t0);
    int64_t nsteps = 1;
    int32_t st = 0;
    while (sign * t0 < sign * tmax) {
        jetcoeffs_generic(pb, order, t, x, params);
        double dt = stepsize(x, order, *o);
        if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
        dt = sign * jl_min(dt, sign * (tmax - t0));
        for (int j = 0; j < dof; j++) x0[j] = horner(x[j], dt);
        t0 += dt;
        t.c[0] = t0;
        for (int j = 0; j < dof; j++) x[j].c[0] = x0[j];
        nsteps += 1;
        if (tv) tv[nsteps - 1] = t0;
        if (xv) for (int j = 0; j < dof; j++) xv[(size_t)j * stride + (nsteps - 1)] = x0[j];
        if (nsteps > o->maxsteps) { st |= TB_ST_MAXSTEPS; break; }
    }
    for (int j = 0; j < dof; j++) xfinal[j] = x0[j];
    *tfinal = t0;
    *nsteps_out = nsteps - 1;
    *status = st;
    return TB_OK;
}

} /* extern "C" */
