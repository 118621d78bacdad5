/*
 * tb_tn_oracle.cpp -- CPU oracle for TaylorN jet transport.  TEST INFRASTRUCTURE ONLY
 * (imported by tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke(); never by the product).
 *
 * Restates the reference's driver with polynomial-valued coefficients, U = TaylorN{Float64}:
 *   taylorinteg!            src/integrator/taylorinteg.jl:329-381 (t0,tmax), 582-642 (trange)
 *   solcoeff! (Taylor1{TaylorN})  src/TaylorIntegration.jl:54-65
 *   _norminf / stepsize on nested coefficients   src/integrator/stepsize.jl:26-42, 58-132
 * The arithmetic is TaylorSeries.jl's (0.22.x, NOT in /root/reference): the Taylor1 recurrences of
 * SURVEY App. B with every coefficient product replaced by a truncated TaylorN product, and the
 * TaylorN functions (/, ^r, sincos) as the same recurrences over homogeneous degree.
 * PARITY UNPINNED at the bit level: the summation order inside TaylorSeries' TaylorN routines is
 * recalled, not read.  Pinned by the reference's own tests at their tolerances
 * (test/jettransport.jl:19-86, 705-805): see tests/test_tn.py.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../include/taylorb200.h"
#include "../taylorintegration.jl_b200/csrc/tb_math.h"
#include "../taylorintegration.jl_b200/csrc/tb_tn_tables.h"

namespace {

struct TN {
    tbtn::Tables T;
    std::vector<int32_t> pos; /* [ncoef*ncoef] flat index of mono[ia]+mono[ib], -1 past the truncation */
    int nc;
    explicit TN(int n, int d) : T(tbtn::make_tables(n, d)), nc(T.ncoef) {
        pos.assign((size_t)nc * nc, -1);
        for (int m = 0; m < nc; m++)
            for (int p = T.pair_off[m]; p < T.pair_off[m + 1]; p++) pos[(size_t)T.pair_a[p] * nc + T.pair_b[p]] = m;
    }
    int lo(int g) const { return T.deg_off[g]; }
    int hi(int g) const { return T.deg_off[g + 1]; }

    /* C[deg ga+gb] += (f*A_ga) x B_gb : HomogeneousPolynomial mul!, scatter form */
    void mulacc_h(double *C, const double *A, int ga, const double *B, int gb, bool hasf, double f) const {
        if (ga + gb > T.d) return;
        for (int ia = lo(ga); ia < hi(ga); ia++) {
            const double ca = hasf ? f * A[ia] : A[ia];
            for (int ib = lo(gb); ib < hi(gb); ib++) C[pos[(size_t)ia * nc + ib]] += ca * B[ib];
        }
    }
    /* truncated TaylorN product, C += (f*A) x B; degrees of A ascending (per slot: ascending ia) */
    void mulacc(double *C, const double *A, const double *B, bool hasf = false, double f = 0.0) const {
        for (int ga = 0; ga <= T.d; ga++)
            for (int gb = 0; ga + gb <= T.d; gb++) mulacc_h(C, A, ga, B, gb, hasf, f);
    }
    /* Q = Nn / B by homogeneous degree: Q_g = (Nn_g - sum_{j<g} Q_j B_{g-j}) / B_0 */
    void div(double *Q, const double *Nn, const double *B) const {
        std::vector<double> tmp(nc, 0.0);
        for (int g = 0; g <= T.d; g++) {
            for (int j = 0; j < g; j++) mulacc_h(tmp.data(), Q, j, B, g - j, false, 0.0);
            for (int m = lo(g); m < hi(g); m++) Q[m] = (Nn[m] - tmp[m]) / B[0];
        }
    }
    /* C = A^r: C_0 = A_0^r, C_g = sum_{j<g} ((r(g-j) - j) C_j) A_{g-j} / (g A_0) */
    void pow(double *C, const double *A, double r) const {
        std::vector<double> tmp(nc, 0.0);
        C[0] = tb_pow(A[0], r);
        for (int g = 1; g <= T.d; g++) {
            for (int j = 0; j < g; j++) mulacc_h(tmp.data(), C, j, A, g - j, true, r * (double)(g - j) - (double)j);
            const double den = (double)g * A[0];
            for (int m = lo(g); m < hi(g); m++) C[m] = tmp[m] / den;
        }
    }
    /* S = sin A, Cc = cos A: S_g = sum_{j=1..g} (j A_j) Cc_{g-j} / g, Cc_g = -sum (j A_j) S_{g-j} / g */
    void sincos(double *S, double *Cc, const double *A) const {
        std::vector<double> ts(nc, 0.0), tc(nc, 0.0);
        tb_sincos(A[0], &S[0], &Cc[0]);
        for (int g = 1; g <= T.d; g++) {
            for (int j = 1; j <= g; j++) {
                mulacc_h(ts.data(), A, j, Cc, g - j, j >= 2, (double)j);
                mulacc_h(tc.data(), A, j, S, g - j, j >= 2, (double)j);
            }
            for (int m = lo(g); m < hi(g); m++) { S[m] = ts[m] / (double)g; Cc[m] = -tc[m] / (double)g; }
        }
    }
    double norminf(const double *A) const {
        double v = 0.0;
        for (int m = 0; m < nc; m++) v = std::fmax(v, std::fabs(A[m]));
        return v;
    }
};

/* ---- Taylor1{TaylorN} series and the tape interpreter ------------------------------------------ */
struct Work {
    const TN &tn;
    const tb_ode_desc_t &d;
    int P, nc;
    std::vector<std::vector<double>> x, dx, tmp; /* each (P+1)*nc, coefficient k at [k*nc] */
    std::vector<double> tser;                    /* the independent variable */
    const double *par = nullptr;
    Work(const TN &t, const tb_ode_desc_t &dd, int order) : tn(t), d(dd), P(order), nc(t.nc) {
        x.assign(d.dof, std::vector<double>((size_t)(P + 1) * nc, 0.0));
        dx.assign(d.dof, std::vector<double>((size_t)(P + 1) * nc, 0.0));
        tmp.assign(d.ntmp > 0 ? d.ntmp : 1, std::vector<double>((size_t)(P + 1) * nc, 0.0));
        tser.assign((size_t)(P + 1) * nc, 0.0);
    }
    double *ser(int32_t o) {
        switch (TB_OPERAND_KIND(o)) {
            case TB_K_X: return x[TB_OPERAND_INDEX(o)].data();
            case TB_K_DX: return dx[TB_OPERAND_INDEX(o)].data();
            case TB_K_TMP: return tmp[TB_OPERAND_INDEX(o)].data();
            case TB_K_TIME: return tser.data();
            default: return nullptr;
        }
    }
    double *co(int32_t o, int k) { return ser(o) + (size_t)k * nc; }
    static bool isnum(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_PAR || k == TB_K_IMM; }
    double num(int32_t o, const tb_op_t &op) const { return TB_OPERAND_KIND(o) == TB_K_PAR ? par[TB_OPERAND_INDEX(o)] : op.imm; }
    void zero(double *p) { std::memset(p, 0, sizeof(double) * nc); }

    /* one tape entry at order k: TaylorSeries.<op>!(res, a, b, k) on Taylor1{TaylorN} */
    int run_op(const tb_op_t &op, int k) {
        const bool sa = isnum(op.a), sb = isnum(op.b);
        double *R = co(op.res, k);
        int opc = op.op;
        if (opc == TB_OP_POW) {
            if (op.imm == 1.0) opc = TB_OP_IDENTITY;
            else if (op.imm == 2.0) opc = TB_OP_SQR;
        }
        switch (opc) {
            case TB_OP_IDENTITY:
                if (sa) { zero(R); if (k == 0) R[0] = num(op.a, op); }
                else std::memcpy(R, co(op.a, k), sizeof(double) * nc);
                break;
            case TB_OP_NEG: { const double *A = co(op.a, k); for (int m = 0; m < nc; m++) R[m] = -A[m]; break; }
            case TB_OP_ADD:
            case TB_OP_SUB: {
                const double sg = opc == TB_OP_ADD ? 1.0 : -1.0;
                if (sa && sb) { zero(R); if (k == 0) R[0] = num(op.a, op) + sg * num(op.b, op); }
                else if (sa) { /* scalar (+|-) series: the scalar touches the constant term of order 0 only */
                    const double *B = co(op.b, k);
                    for (int m = 0; m < nc; m++) R[m] = opc == TB_OP_ADD ? B[m] : -B[m];
                    if (k == 0) R[0] = opc == TB_OP_ADD ? num(op.a, op) + B[0] : num(op.a, op) - B[0];
                } else if (sb) {
                    const double *A = co(op.a, k);
                    std::memcpy(R, A, sizeof(double) * nc);
                    if (k == 0) R[0] = opc == TB_OP_ADD ? A[0] + num(op.b, op) : A[0] - num(op.b, op);
                } else {
                    const double *A = co(op.a, k), *B = co(op.b, k);
                    for (int m = 0; m < nc; m++) R[m] = opc == TB_OP_ADD ? A[m] + B[m] : A[m] - B[m];
                }
                break;
            }
            case TB_OP_MUL:
                if (sa && sb) { zero(R); if (k == 0) R[0] = num(op.a, op) * num(op.b, op); }
                else if (sa || sb) {
                    const double s = sa ? num(op.a, op) : num(op.b, op);
                    const double *A = co(sa ? op.b : op.a, k);
                    for (int m = 0; m < nc; m++) R[m] = s * A[m];
                } else {
                    zero(R);
                    for (int i = 0; i <= k; i++) tn.mulacc(R, co(op.a, i), co(op.b, k - i));
                }
                break;
            case TB_OP_SQR: {
                zero(R);
                if (k == 0) { tn.mulacc(R, co(op.a, 0), co(op.a, 0)); break; }
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                for (int i = 0; i <= kend; i++) tn.mulacc(R, co(op.a, i), co(op.a, k - i));
                for (int m = 0; m < nc; m++) R[m] = 2.0 * R[m];
                if (!kodd) tn.mulacc(R, co(op.a, k / 2), co(op.a, k / 2));
                break;
            }
            case TB_OP_DIV:
                if (sb) { /* series / scalar */
                    if (sa) { zero(R); if (k == 0) R[0] = num(op.a, op) / num(op.b, op); break; }
                    const double s = num(op.b, op);
                    const double *A = co(op.a, k);
                    for (int m = 0; m < nc; m++) R[m] = A[m] / s;
                } else {
                    std::vector<double> t(nc, 0.0), nn(nc, 0.0);
                    for (int i = 0; i < k; i++) tn.mulacc(t.data(), co(op.res, i), co(op.b, k - i));
                    if (sa) { if (k == 0) nn[0] = num(op.a, op); }
                    else std::memcpy(nn.data(), co(op.a, k), sizeof(double) * nc);
                    for (int m = 0; m < nc; m++) nn[m] = nn[m] - t[m];
                    tn.div(R, nn.data(), co(op.b, 0));
                }
                break;
            case TB_OP_POW:
                if (sa) return TB_ERR_UNSUPPORTED;
                if (k == 0) tn.pow(R, co(op.a, 0), op.imm);
                else {
                    std::vector<double> t(nc, 0.0), den(nc);
                    for (int i = 0; i < k; i++)
                        tn.mulacc(t.data(), co(op.res, i), co(op.a, k - i), true, op.imm * (double)(k - i) - (double)i);
                    const double *A0 = co(op.a, 0);
                    for (int m = 0; m < nc; m++) den[m] = (double)k * A0[m];
                    tn.div(R, t.data(), den.data());
                }
                break;
            case TB_OP_SINCOS: {
                if (sa) return TB_ERR_UNSUPPORTED;
                double *Cc = co(op.aux, k);
                if (k == 0) { tn.sincos(R, Cc, co(op.a, 0)); break; }
                std::vector<double> ts(nc, 0.0), tc(nc, 0.0);
                for (int i = 1; i <= k; i++) {
                    tn.mulacc(ts.data(), co(op.a, i), co(op.aux, k - i), i >= 2, (double)i);
                    tn.mulacc(tc.data(), co(op.a, i), co(op.res, k - i), i >= 2, (double)i);
                }
                for (int m = 0; m < nc; m++) { R[m] = ts[m] / (double)k; Cc[m] = -tc[m] / (double)k; }
                break;
            }
            default: return TB_ERR_UNSUPPORTED;
        }
        return 0;
    }
    /* jetcoeffs!: parse_eqs.jl:259-264 + solcoeff! (TaylorIntegration.jl:54-65) */
    int jetcoeffs(double t0) {
        std::fill(tser.begin(), tser.end(), 0.0);
        tser[0] = t0;
        if (P >= 1) tser[(size_t)nc] = 1.0;
        for (int k = 0; k < P; k++) {
            for (int j = 0; j < d.nops; j++) { int rc = run_op(d.ops[j], k); if (rc) return rc; }
            for (int c = 0; c < d.dof; c++) {
                const double *D = dx[c].data() + (size_t)k * nc;
                double *X = x[c].data() + (size_t)(k + 1) * nc;
                for (int m = 0; m < nc; m++) X[m] = D[m] / (double)(k + 1);
            }
        }
        return 0;
    }
    /* stepsize(q::AbstractArray{Taylor1{TaylorN}}): stepsize.jl:58-99 with _norminf 26-42 */
    double stepsize(double abstol, double reltol) {
        double h = std::numeric_limits<double>::infinity();
        for (int c = 0; c < d.dof; c++) {
            const double x0norm = tn.norminf(x[c].data());
            for (int k = P - 1; k <= P; k++) {
                const double a = tn.norminf(x[c].data() + (size_t)k * nc);
                if (a == 0.0) continue;
                const double rel = reltol * x0norm;
                const double eps = abstol >= rel ? abstol : rel;
                h = std::min(h, tb_pow(eps / a, 1.0 / k));
            }
        }
        if (std::isinf(h)) { /* _second_stepsize: stepsize.jl:119-132 */
            h = 0.0;
            for (int c = 0; c < d.dof; c++) {
                bool allzero = true;
                for (size_t j = 0; j < x[c].size(); j++) if (x[c][j] != 0.0) { allzero = false; break; }
                double hc = std::numeric_limits<double>::infinity();
                if (!allzero) {
                    hc = 0.0;
                    for (int k = 1; k <= P - 2; k++) {
                        const double a = tn.norminf(x[c].data() + (size_t)k * nc);
                        if (a == 0.0) continue;
                        hc = std::max(hc, tb_pow(1.0 / a, 1.0 / k));
                    }
                }
                h = std::max(h, hc);
            }
        }
        return h;
    }
    /* evaluate(x[c], dt): Horner on every polynomial coefficient */
    void horner(int c, double dt, double *out) const {
        const double *X = x[c].data();
        for (int m = 0; m < nc; m++) {
            double s = X[(size_t)P * nc + m];
            for (int k = P - 1; k >= 0; k--) s = s * dt + X[(size_t)k * nc + m];
            out[m] = s;
        }
    }
};

} // namespace

extern "C" {

/* monomial exponents [ncoef x nvars] in the flat order; returns ncoef */
int tbo_tn_monomials(int nvars, int varorder, int32_t *expo) {
    tbtn::Tables T = tbtn::make_tables(nvars, varorder);
    if (expo) for (size_t i = 0; i < T.expo.size(); i++) expo[i] = T.expo[i];
    return T.ncoef;
}

/* Ensemble of jet-transport integrations.  q0 [dof x ncoef x N] (trajectory index fastest), outputs
 * xfinal [dof x ncoef x N], tfinal/nsteps/status [N], xv_trange [dof x ncoef x nt x N] (NaN-prefilled).
 * trange == NULL: (t0, tmax) driver, taylorinteg.jl:329-381; else the trange driver, 582-642.          */
int tbo_tn_integrate(const tb_ode_desc_t *desc, int order, int nvars, int varorder, int64_t N, const double *q0,
                     const double *params, double t0_, double tmax_, const double *trange, int64_t nt,
                     const tb_opts_t *o, double *xfinal, double *tfinal, int64_t *nsteps_out, int32_t *status,
                     double *xv_trange) {
    TN tn(nvars, varorder);
    const int nc = tn.nc, dof = desc->dof;
    const double nan = std::numeric_limits<double>::quiet_NaN();
    std::vector<double> par(desc->npar > 0 ? desc->npar : 1), xn((size_t)dof * nc);
    for (int64_t i = 0; i < N; i++) {
        Work w(tn, *desc, order);
        for (int j = 0; j < desc->npar; j++) par[j] = o->params_per_traj ? params[(int64_t)j * N + i] : params[j];
        w.par = par.data();
        for (int c = 0; c < dof; c++)
            for (int m = 0; m < nc; m++) w.x[c][m] = q0[((int64_t)c * nc + m) * N + i];
        double t0 = trange ? trange[0] : t0_;
        const double tmax = trange ? trange[nt - 1] : tmax_;
        const double sign = std::copysign(1.0, tmax - t0);
        int64_t nsteps = 1, iter = 2;
        double t1 = trange ? trange[1] : 0.0;
        int st = 0;
        if (trange && xv_trange)
            for (int c = 0; c < dof; c++)
                for (int m = 0; m < nc; m++) {
                    xv_trange[(((int64_t)c * nc + m) * nt) * N + i] = w.x[c][m];
                    for (int64_t j = 1; j < nt; j++) xv_trange[(((int64_t)c * nc + m) * nt + j) * N + i] = nan;
                }
        while (sign * t0 < sign * tmax) {
            int rc = w.jetcoeffs(t0);
            if (rc) return rc;
            double dt = w.stepsize(o->abstol, o->reltol);
            dt = dt < o->minstepsize ? o->minstepsize : (dt > o->maxstepsize ? o->maxstepsize : dt);
            if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
            const double rem = tmax - t0;
            dt = sign * std::min(dt, sign * rem);
            for (int c = 0; c < dof; c++) w.horner(c, dt, xn.data() + (size_t)c * nc);
            if (trange) {
                const double tnext = t0 + dt;
                std::vector<double> xo(nc);
                while (sign * t1 < sign * tnext) {
                    for (int c = 0; c < dof; c++) {
                        w.horner(c, t1 - t0, xo.data());
                        if (xv_trange) for (int m = 0; m < nc; m++) xv_trange[(((int64_t)c * nc + m) * nt + (iter - 1)) * N + i] = xo[m];
                    }
                    iter += 1;
                    t1 = trange[iter - 1];
                }
                for (int c = 0; c < dof; c++) std::memcpy(w.x[c].data(), xn.data() + (size_t)c * nc, sizeof(double) * nc);
                if (dt == rem) {
                    if (xv_trange)
                        for (int c = 0; c < dof; c++)
                            for (int m = 0; m < nc; m++) xv_trange[(((int64_t)c * nc + m) * nt + (iter - 1)) * N + i] = w.x[c][m];
                    t0 = tnext;
                    break;
                }
                t0 = tnext;
                nsteps += 1;
            } else {
                for (int c = 0; c < dof; c++) std::memcpy(w.x[c].data(), xn.data() + (size_t)c * nc, sizeof(double) * nc);
                t0 = t0 + dt;
                nsteps += 1;
            }
            if (nsteps > o->maxsteps) { st |= TB_ST_MAXSTEPS; break; }
            bool fin = true;
            for (int c = 0; c < dof; c++) for (int m = 0; m < nc; m++) fin = fin && std::isfinite(w.x[c][m]);
            if (!fin) { st |= TB_ST_NONFINITE; break; }
        }
        if (xfinal) for (int c = 0; c < dof; c++) for (int m = 0; m < nc; m++) xfinal[((int64_t)c * nc + m) * N + i] = w.x[c][m];
        if (tfinal) tfinal[i] = t0;
        if (nsteps_out) nsteps_out[i] = nsteps - 1;
        if (status) status[i] = st;
    }
    return 0;
}

} /* extern "C" */
