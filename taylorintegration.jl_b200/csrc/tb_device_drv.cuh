/*
 * tb_device_drv.cuh -- second half of the NVRTC translation unit: the fixed
 * (tape-independent) part of the fused kernel.  One thread integrates one
 * trajectory from t0 to tmax without leaving the SM: jet coefficients
 * (generated code), step size, Horner evaluation, event location, Lyapunov
 * QR, output stores, all in one launch.
 *
 * Every function cites the reference routine it restates (paths relative to
 * the TaylorIntegration.jl tree).  In parity mode each floating-point
 * operation below is the same IEEE operation, in the same order, as in
 * oracle/tb_oracle.cpp.
 */
#ifndef TB_DEVICE_DRV_CUH
#define TB_DEVICE_DRV_CUH

#ifndef TB_VARIANT
#define TB_VARIANT 1
#endif
#ifndef TB_BAR
#define TB_BAR 0
#endif
/* TB_NQ: components of the ODE proper (step-size control, finiteness, outputs).  It equals TB_DOF except in the
 * column-parallel Lyapunov kernel (TB_KIND == 3), whose jets cover [x; v] = 2*TB_NQ components.              */
#ifndef TB_NQ
#define TB_NQ TB_DOF
#endif
#ifndef TB_LYAP_TRANGE
#define TB_LYAP_TRANGE 0
#endif
#ifndef TB_FINAL_ONLY
#define TB_FINAL_ONLY 0
#endif

/* _stepsize: src/integrator/stepsize.jl:107-111   (eps/a)^(1/k).  Not inlined: the
 * double-double pow is ~300 instructions and is called 2*dof times per step.    */
__device__ __noinline__ double tb_stepsize_k(double a, double eps, int k) {
    double aux = D_(eps, a);
    double kinv = D_(1.0, (double)k);
    return tb_pow(aux, kinv);
}

#ifndef TB_STEPSIZE_PAIR
#define TB_STEPSIZE_PAIR 0
#endif
#if TB_STEPSIZE_PAIR
/* (eps/a1)^(1/(P-1)) and (eps/a2)^(1/P) together: same bits as two tb_stepsize_k calls (1/k is the same correctly
 * rounded quotient, folded at compile time).  Measured: -7 % on Kepler-28 parity (232 registers of jets live across
 * the call, the wider callee spills) and -3 % on pendulum-25: the other warps already cover the latency.  Opt-in. */
__device__ __noinline__ tb_pair tb_stepsize_pair(double a1, double a2, double eps) {
    const double x1 = D_(eps, a1), x2 = D_(eps, a2);
    return tb_pow_pair(x1, 1.0 / (double)(TB_P - 1), x2, 1.0 / (double)TB_P);
}
#endif

#if TB_VARIANT == 1
/* evaluate(a::Taylor1, dt): Horner, s = a[P]; s = s*dt + a[k]  (TaylorSeries) */
__device__ __forceinline__ double tb_horner(const double *__restrict__ smt, int slot, double dt) {
    double s = S(slot, TB_P);
#pragma unroll 4
    for (int k = TB_P - 1; k >= 0; k--) s = A_(M_(s, dt), S(slot, k));
    return s;
}

/* stepsize(x::Taylor1, abs, rel): stepsize.jl:58-77 */
__device__ __forceinline__ double tb_stepsize1(const double *__restrict__ smt, int slot, double abseps, double releps) {
    double x0norm = fabs(S(slot, 0));
    double h = TB_INF;
#pragma unroll
    for (int k = TB_P - 1; k <= TB_P; k++) {
        double aux = fabs(S(slot, k));
        if (aux == 0.0) continue;
        double rel = M_(releps, x0norm);
        double eps = (abseps >= rel) ? abseps : rel;
        double h1 = tb_stepsize_k(aux, eps, k);
        h = tb_jmin(h, h1);
    }
    return h;
}

/* _second_stepsize: stepsize.jl:119-132 */
__device__ __noinline__ double tb_second_stepsize(const double *__restrict__ smt, int slot) {
    bool allzero = true;
    for (int k = 0; k <= TB_P; k++)
        if (S(slot, k) != 0.0) { allzero = false; break; }
    if (allzero) return TB_INF;
    double h = 0.0;
    for (int k = 1; k <= TB_P - 2; k++) {
        double aux = fabs(S(slot, k));
        if (aux == 0.0) continue;
        h = tb_jmax(h, tb_stepsize_k(aux, 1.0, k));
    }
    return h;
}

/* stepsize(q::AbstractArray{Taylor1}, abs, rel): stepsize.jl:79-99, then the
 * clamp of taylorstep! (src/integrator/taylorstep.jl:79).  Only the first
 * TB_DOF jets enter (lyap_taylorstep!, lyapunovspectrum.jl:270).             */
__device__ __forceinline__ double tb_stepsize(const double *__restrict__ smt, const TbArgs &a) {
    double h = TB_INF;
#pragma unroll
    for (int i = 0; i < TB_DOF; i++) h = tb_jmin(h, tb_stepsize1(smt, i, a.abstol, a.reltol));
    if (h == TB_INF) {
        h = 0.0;
        for (int i = 0; i < TB_DOF; i++) h = tb_jmax(h, tb_second_stepsize(smt, i));
    }
    /* clamp(h, lo, hi) */
    h = (h < a.minstep) ? a.minstep : ((h > a.maxstep) ? a.maxstep : h);
    return h;
}

/* jetcoeffs!(::Val{f}, t, x, dx, params, rv): skeleton src/parse_eqs.jl:259-264 */
__device__ __forceinline__ void tb_jetcoeffs(double *__restrict__ smt, const double *par, double t0) {
    tb_rhs_k0(smt, par, t0);
#pragma unroll 1
    for (int k = 1; k < TB_P; k++) tb_rhs_k(smt, par, t0, k);
}

#endif /* TB_VARIANT == 1 */

#if TB_VARIANT == 2
/* ---- variant 2: jets in registers / shared memory / derived, static indices ----
 * TB_X(c, k) is coefficient k of state jet c; every loop over (c, k) below has
 * constant bounds and is fully unrolled, so each access resolves at compile time
 * to a register, one LDS, or a scaled dx coefficient.                            */
#define TB_X(c, k) tb_u_x(J, smt, x0, (c), (k))

__device__ __forceinline__ double tb_u_horner(const TbJets &J, const double *__restrict__ smt,
                                              const double *__restrict__ x0, const int c, const double dt) {
    double s = TB_X(c, TB_P);
#pragma unroll
    for (int k = TB_P - 1; k >= 0; k--) s = A_(M_(s, dt), TB_X(c, k));
    return s;
}

/* _second_stepsize on a local copy of one jet (cold path) */
__device__ __noinline__ double tb_second_stepsize_arr(const double *x) {
    bool allzero = true;
    for (int k = 0; k <= TB_P; k++)
        if (x[k] != 0.0) { allzero = false; break; }
    if (allzero) return TB_INF;
    double h = 0.0;
    for (int k = 1; k <= TB_P - 2; k++) {
        double aux = fabs(x[k]);
        if (aux == 0.0) continue;
        h = tb_jmax(h, tb_stepsize_k(aux, 1.0, k));
    }
    return h;
}

/* stepsize + clamp (stepsize.jl:58-99, taylorstep.jl:79).  Fast mode with
 * reltol == 0: min over components of (eps/|x_k|)^(1/k) = (eps/max_c|x_k|)^(1/k),
 * so two logarithms and one exponential per step instead of 2*dof pow calls.     */
__device__ __forceinline__ double tb_u_stepsize(const TbJets &J, const double *__restrict__ smt,
                                                const double *__restrict__ x0, const TbArgs &a) {
    double h = TB_INF;
#if TB_MODE == 1
    if (a.reltol == 0.0) {
        double m1 = 0.0, m2 = 0.0;
#pragma unroll
        for (int c = 0; c < TB_NQ; c++) {
            m1 = fmax(m1, fabs(TB_X(c, TB_P - 1)));
            m2 = fmax(m2, fabs(TB_X(c, TB_P)));
        }
        const double le = log(a.abstol);
        double l = TB_INF;
        if (m1 != 0.0) l = fmin(l, (le - log(m1)) * (1.0 / (double)(TB_P - 1)));
        if (m2 != 0.0) l = fmin(l, (le - log(m2)) * (1.0 / (double)TB_P));
        if (l != TB_INF) h = exp(l);
    } else
#else
    if (a.reltol == 0.0) {
        /* Parity mode, reltol == 0: min_c pow(eps/|x_c[k]|, 1/k) is attained at the largest
         * |x_c[k]| (IEEE division is monotone; tb_pow errs by < 1 ulp, so two arguments that
         * differ by more than 2^-40 relative cannot swap order after the 1/k-th power).  Only
         * components within 2^-40 of the maximum are evaluated as well: same bits as the
         * reference's loop over all components (stepsize.jl:83-86), 2 pow calls instead of 2*dof.
         * Julia's min PROPAGATES NaN: one NaN |x_c[k]| makes h NaN (then dt, the state and t0, and the
         * driver loop ends); fmax drops NaNs, so they are tracked separately (`anynan`).          */
        bool anynan = false;
#if TB_STEPSIZE_PAIR
        /* TB_CF_PAIR_POW: the two powers in ONE straight-line body (tb_pow_pair, same bits) so that their dependent
         * chains overlap; measured slower than two calls on both benchmark kernels, kept as an option            */
        double mk2[2];
#pragma unroll
        for (int j = 0; j < 2; j++) {
            mk2[j] = 0.0;
#pragma unroll
            for (int c = 0; c < TB_NQ; c++) mk2[j] = fmax(mk2[j], fabs(TB_X(c, TB_P - 1 + j)));
        }
        const bool both = mk2[0] != 0.0 && mk2[1] != 0.0;
        tb_pair hp;
        hp.a = 0.0; hp.b = 0.0;
        if (both) hp = tb_stepsize_pair(mk2[0], mk2[1], a.abstol);
#endif
        double acs[2][TB_NQ]; /* every coefficient is read before the first data-dependent branch (tensor-memory loads are warp-collective) */
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int c = 0; c < TB_NQ; c++) acs[j][c] = fabs(TB_X(c, TB_P - 1 + j));
        /* Which of the two candidates (eps/m1)^(1/(P-1)), (eps/m2)^(1/P) is the minimum is decided from single-precision
         * logarithms (exponent + MUFU.LG2 of the mantissa, absolute error ~1e-6 in log2 h) whenever they differ by more
         * than 1e-4 in log2 -- tb_pow errs by < 1 ulp, so the order of the two double-double powers cannot differ -- and
         * only that one is evaluated: one ~300-instruction tb_pow per step instead of two, same bits.                */
        int skip_k = -1;
#if !TB_STEPSIZE_PAIR
        {
            double mm[2] = {0.0, 0.0};
#pragma unroll
            for (int j = 0; j < 2; j++)
#pragma unroll
                for (int c = 0; c < TB_NQ; c++) mm[j] = fmax(mm[j], acs[j][c]);
            if (mm[0] > 0.0 && mm[1] > 0.0 && mm[0] < TB_INF && mm[1] < TB_INF && a.abstol > 0.0) {
                float l2[3];
                const double v[3] = {a.abstol, mm[0], mm[1]};
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    const int hi = __double2hiint(v[j]);
                    const int ex = ((hi >> 20) & 0x7ff) - 1023; /* subnormals land far off: harmless, only a hint */
                    const double mant = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(v[j]));
                    l2[j] = (float)ex + __log2f((float)mant);
                }
                const float e1 = (l2[0] - l2[1]) * (1.0f / (float)(TB_P - 1)), e2 = (l2[0] - l2[2]) * (1.0f / (float)TB_P);
                /* eps/m must be a normal number as well: a subnormal (or zero) quotient is not what the logarithms describe */
                const bool normal = fabsf(l2[0] - l2[1]) < 1000.0f && fabsf(l2[0] - l2[2]) < 1000.0f &&
                                    ((__double2hiint(mm[0]) >> 20) & 0x7ff) != 0 && ((__double2hiint(mm[1]) >> 20) & 0x7ff) != 0 &&
                                    ((__double2hiint(a.abstol) >> 20) & 0x7ff) != 0;
                if (normal && e1 - e2 > 1e-4f) skip_k = TB_P - 1;      /* the k = P candidate is the smaller one */
                else if (normal && e2 - e1 > 1e-4f) skip_k = TB_P;
            }
        }
#endif
#pragma unroll
        for (int k = TB_P - 1; k <= TB_P; k++) {
            double m = 0.0, ac[TB_NQ];
#pragma unroll
            for (int c = 0; c < TB_NQ; c++) { ac[c] = acs[k - (TB_P - 1)][c]; m = fmax(m, ac[c]); anynan = anynan || (ac[c] != ac[c]); }
            if (m != 0.0 && k != skip_k) {
#if TB_STEPSIZE_PAIR
                double hk = both ? (k == TB_P ? hp.b : hp.a) : tb_stepsize_k(m, a.abstol, k);
#else
                double hk = tb_stepsize_k(m, a.abstol, k);
#endif
                const double thr = M_(m, 0.99999999999909051);
#pragma unroll
                for (int c = 0; c < TB_NQ; c++)
                    if (ac[c] != m && ac[c] > thr) hk = fmin(hk, tb_stepsize_k(ac[c], a.abstol, k));
                h = fmin(h, hk);
            }
        }
        if (anynan) h = TB_NAN;
    } else
#endif
    {
        double xa[3][TB_NQ];
#pragma unroll
        for (int c = 0; c < TB_NQ; c++) { xa[0][c] = fabs(TB_X(c, 0)); xa[1][c] = fabs(TB_X(c, TB_P - 1)); xa[2][c] = fabs(TB_X(c, TB_P)); }
#pragma unroll
        for (int c = 0; c < TB_NQ; c++) {
            const double x0norm = xa[0][c];
#pragma unroll
            for (int k = TB_P - 1; k <= TB_P; k++) {
                const double aux = xa[1 + k - (TB_P - 1)][c];
                if (aux == 0.0) continue;
                const double rel = M_(a.reltol, x0norm);
                const double eps = (a.abstol >= rel) ? a.abstol : rel;
                h = tb_jmin(h, tb_stepsize_k(aux, eps, k));
            }
        }
    }
    const bool need2 = h == TB_INF; /* all last coefficients vanish: second step-size control */
#if TB_TMEM
    if (__any_sync(0xffffffffu, need2)) { /* the coefficient reads below may be warp-collective: the whole warp takes the branch */
#else
    if (need2) {
#endif
        double h2 = 0.0;
        double buf[TB_P + 1];
#pragma unroll
        for (int c = 0; c < TB_NQ; c++) {
#pragma unroll
            for (int k = 0; k <= TB_P; k++) buf[k] = TB_X(c, k);
            if (need2) h2 = tb_jmax(h2, tb_second_stepsize_arr(buf));
        }
        if (need2) h = h2;
#if TB_TMEM
        __syncwarp();
#endif
    }
    h = (h < a.minstep) ? a.minstep : ((h > a.maxstep) ? a.maxstep : h);
    return h;
}
#endif /* TB_VARIANT == 2 */

#if TB_KIND == 2 || TB_KIND == 3
/* modifiedGS!(A, Q, R): src/lyapunovspectrum.jl:118-146, column-major n x n;
 * dot and norm as left-to-right sums (oracle: modifiedGS).                    */
__device__ __forceinline__ void tb_mgs(const double *A, double *Q, double *Rd) {
    double v[TB_NQ];
#pragma unroll
    for (int j = 0; j < TB_NQ; j++) {
#pragma unroll
        for (int r = 0; r < TB_NQ; r++) v[r] = A[r + j * TB_NQ];
#pragma unroll
        for (int i = 0; i < TB_NQ; i++) {
            if (i < j) {
                double dot = M_(Q[0 + i * TB_NQ], v[0]);
#pragma unroll
                for (int r = 1; r < TB_NQ; r++) dot = A_(dot, M_(Q[r + i * TB_NQ], v[r]));
#pragma unroll
                for (int r = 0; r < TB_NQ; r++) v[r] = B_(v[r], M_(dot, Q[r + i * TB_NQ]));
            }
        }
        double ss = M_(v[0], v[0]);
#pragma unroll
        for (int r = 1; r < TB_NQ; r++) ss = A_(ss, M_(v[r], v[r]));
        double nrm = Q_(ss);
        Rd[j] = nrm;
#pragma unroll
        for (int r = 0; r < TB_NQ; r++) Q[r + j * TB_NQ] = D_(v[r], nrm);
    }
}
#endif

#if TB_KIND == 1
/* coefficient k of derivative(g, eo) (repeated d[k] = (k+1) a[k+1], TaylorSeries) */
__device__ __forceinline__ double tb_gder(const double *__restrict__ smt, int eo, int k) {
    if (k + eo > TB_P) return 0.0;
    double v = TB_EV_G(k + eo);
    for (int r = eo; r >= 1; r--) v = M_((double)(k + r), v);
    return v;
}
/* evaluate!(g_dg, dt): G = derivative(g, eo), dG = derivative(G), both Horner */
__device__ __forceinline__ void tb_eval_g_dg(const double *__restrict__ smt, int eo, double dt, double *gv, double *dgv) {
    if (eo == 0) {
        /* the usual case (eventorder = 0, G = g): same operations with static indices, each coefficient read once */
        double gk1 = TB_EV_G(TB_P), s0 = gk1, ds0 = 0.0;
#pragma unroll
        for (int k = TB_P - 1; k >= 0; k--) {
            ds0 = A_(M_(ds0, dt), M_((double)(k + 1), gk1));
            gk1 = TB_EV_G(k);
            s0 = A_(M_(s0, dt), gk1);
        }
        *gv = s0;
        *dgv = ds0;
        return;
    }
    double s = tb_gder(smt, eo, TB_P), ds = 0.0; /* dG[P] = 0 */
    for (int k = TB_P - 1; k >= 0; k--) {
        double Gk1 = tb_gder(smt, eo, k + 1);
        ds = A_(M_(ds, dt), M_((double)(k + 1), Gk1));
        s = A_(M_(s, dt), tb_gder(smt, eo, k));
    }
    *gv = s;
    *dgv = ds;
}
#if TB_VARIANT == 1
/* g(dx, x, params, t) on the current jets: returns the Bool of the tuple */
__device__ __forceinline__ bool tb_eval_event(double *__restrict__ smt, const double *par, double t0) {
#if TB_EV_NOPS > 0
    tb_ev_k0(smt, par, t0);
#pragma unroll 1
    for (int k = 1; k <= TB_P; k++) tb_ev_k(smt, par, t0, k);
#endif
    return TB_EV_GATE_EXPR;
}
#endif
#endif

/* ---- one trajectory ------------------------------------------------------- */
struct TbTraj {
    long long i;      /* trajectory index */
    double t0;        /* current time */
    double x0[TB_NX]; /* current state */
    double par[TB_NPAR > 0 ? TB_NPAR : 1];
    long long nsteps; /* 1-based like the reference */
    int st;
    double sign;
    double tmax;
#if TB_KIND == 1
    double g_old, dt, dt_old;
    bool f_old;
    int nevents; /* 1-based */
#endif
#if TB_KIND == 2
    double lsum[TB_DOF], lam[TB_DOF], t00;
#endif
#if TB_KIND == 3
    double lsum, lam, t00; /* this lane's exponent (column `lit` of the fundamental matrix) */
    int lit;               /* lane in team = column index */
    int leader;            /* warp lane of the team's first member */
    unsigned tmask;        /* the team's lanes */
#endif
    /* trange variant */
    long long iter;
    double t1;
};

__device__ __forceinline__ bool tb_isfinite_state(const double *x0) {
    bool fin = true;
#pragma unroll
    for (int c = 0; c < TB_NQ; c++) fin = fin && isfinite(x0[c]);
    return fin;
}

/* update_cache!: src/integrator/cache.jl:435-442 */
__device__ __forceinline__ void tb_update_cache(double *__restrict__ smt, const double *x0) {
#if TB_VARIANT == 1
#pragma unroll
    for (int c = 0; c < TB_NX; c++) S(c, 0) = x0[c];
#endif /* variant 2: tb_u_jets reads x0 directly */
}

#if TB_KIND != 3
/* returns true when the trajectory needs no step at all */
__device__ __forceinline__ bool tb_traj_init(TbTraj &T, double *__restrict__ smt, const TbArgs &a, long long i) {
    T.i = i;
    const long long N = a.N;
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) T.x0[c] = a.q0[(long long)c * N + i];
#if TB_KIND == 2
    /* init_cache_lyap: state = [q; vec(I)]  src/integrator/cache.jl:320-369 */
#pragma unroll
    for (int c = 0; c < TB_DOF * TB_DOF; c++) T.x0[TB_DOF + c] = ((c % TB_DOF) == (c / TB_DOF)) ? 1.0 : 0.0;
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) { T.lsum[c] = 0.0; T.lam[c] = 0.0; }
    if (a.lyap_state0) { /* continue from state: frame Q and running sums of the previous call (taylorb200.h) */
#pragma unroll
        for (int c = 0; c < TB_DOF * TB_DOF; c++) T.x0[TB_DOF + c] = a.lyap_state0[(long long)c * N + i];
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) T.lsum[c] = a.lyap_state0[(long long)(TB_DOF * TB_DOF + c) * N + i];
    }
#endif
#pragma unroll
    for (int j = 0; j < TB_NPAR; j++) T.par[j] = a.params_per_traj ? a.params[(long long)j * N + i] : a.params[j];
    T.nsteps = 1;
    T.st = 0;
    T.iter = 2;
    if (a.trange) {
        T.t0 = a.trange[0];
        T.t1 = a.trange[1];
        T.tmax = a.trange[a.nt - 1];
        /* NaN prefill (src/integrator/cache.jl:313-316) + first column */
        if (a.xv_trange)
            for (int c = 0; c < TB_DOF; c++) {
                a.xv_trange[((long long)c * a.nt) * N + i] = T.x0[c];
                for (long long j = 1; j < a.nt; j++) a.xv_trange[((long long)c * a.nt + j) * N + i] = TB_NAN;
            }
    } else {
        T.t0 = a.t0;
        T.tmax = a.tmax;
        T.t1 = 0.0;
    }
    T.sign = copysign(1.0, B_(T.tmax, T.t0));
    tb_update_cache(smt, T.x0);
    if (a.tv) a.tv[i] = T.t0;
    if (a.xv)
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) a.xv[((long long)c * a.stride) * N + i] = T.x0[c];
#if TB_KIND == 2
    T.t00 = a.lyap_state0 ? a.lyap_state0[(long long)(TB_DOF * TB_DOF + TB_DOF) * N + i] : T.t0;
    if (a.lambda_v)
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) a.lambda_v[((long long)c * a.stride) * N + i] = 0.0;
#endif
#if TB_KIND == 1
    /* rootfinding.jl:377-378: g on the initial jets, which the parsed-method
     * probe (jetcoeffs.jl:181-182 via cache.jl:266) has already filled.       */
#if TB_VARIANT == 1
    tb_jetcoeffs(smt, T.par, T.t0);
    T.f_old = tb_eval_event(smt, T.par, T.t0);
    T.g_old = TB_EV_G(a.eventorder);
#else
    if (a.eventorder == 0) {
        /* g and the gate are state components: their order-0 coefficients are the state itself, so a lane that
         * refills needs no jet evaluation (which would stall the other warps of a step-lockstep block)          */
        T.f_old = TB_EV_GATE0(T.x0);
        T.g_old = TB_EV_G0(T.x0);
    } else { /* variant 2: g is a state component (generated accessors), jets in registers / shared memory */
        TbJets J;
        J.tm = 0u;
        const double *x0 = T.x0;
        tb_u_jets(J, smt, 0u, T.par, T.t0, x0);
        T.f_old = TB_EV_GATE_EXPR;
        T.g_old = tb_u_gcoef(J, smt, x0, a.eventorder);
    }
#endif
    T.dt = 0.0;
    T.dt_old = 0.0;
    T.nevents = 1;
#endif
    return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
}

#endif /* TB_KIND != 3 */

/* variant-independent spelling of "the jets of the current step" */
#if TB_VARIANT == 1
#define TB_JETS_DECL
#define TB_JETS_ARGS smt
#define TB_JETS_PARAMS const double *__restrict__ smt
#define TB_COMPUTE_JETS(T) tb_jetcoeffs(smt, (T).par, (T).t0)
#define TB_STEPSIZE(a) tb_stepsize(smt, (a))
#define TB_HORNER(c, dt) tb_horner(smt, (c), (dt))
#define TB_XCOEF(c, k) S((c), (k))
#define TB_UNROLL_V2
#else
#define TB_JETS_DECL TbJets J; J.tm = tb_tm; const double *x0 = T.x0;
#define TB_JETS_ARGS J, smt, x0
#define TB_JETS_PARAMS const TbJets &J, const double *__restrict__ smt, const double *__restrict__ x0
#define TB_COMPUTE_JETS(T) tb_u_jets(J, smt, tb_tm, (T).par, (T).t0, x0)
#define TB_STEPSIZE(a) tb_u_stepsize(J, smt, x0, (a))
#define TB_HORNER(c, dt) tb_u_horner(J, smt, x0, (c), (dt))
#define TB_XCOEF(c, k) TB_X((c), (k))
#define TB_UNROLL_V2 _Pragma("unroll")
#endif

#if TB_KIND != 3
__device__ __forceinline__ void tb_store_step(const TbTraj &T, const double *__restrict__ smt, const TbArgs &a) {
    /* called after nsteps += 1: column nsteps-1 */
    const long long s = T.nsteps - 1, N = a.N;
    if (s < a.stride) {
        if (a.tv) a.tv[s * N + T.i] = T.t0;
        if (a.xv)
#pragma unroll
            for (int c = 0; c < TB_DOF; c++) a.xv[((long long)c * a.stride + s) * N + T.i] = T.x0[c];
    }
}
/* set_psol!: src/integrator/taylorinteg.jl:176-186 (polynomial of the step just taken) */
__device__ __forceinline__ void tb_store_psol(const TbTraj &T, TB_JETS_PARAMS, const TbArgs &a) {
    const long long s = T.nsteps - 1, N = a.N;
    if (a.psol && s < a.stride) {
        TB_UNROLL_V2
        for (int c = 0; c < TB_DOF; c++) {
            TB_UNROLL_V2
            for (int k = 0; k <= TB_P; k++)
                a.psol[(((long long)c * TB_NP1 + k) * a.stride + s) * N + T.i] = TB_XCOEF(c, k);
        }
    }
}

/* One iteration of the driver loop; returns true when the trajectory is done. */
__device__ __forceinline__ bool tb_traj_step(TbTraj &T, double *__restrict__ smt, const unsigned tb_tm, const TbArgs &a, const bool live) {
#if TB_KIND == 0
    /* taylorinteg!: src/integrator/taylorinteg.jl:356-378 (t0,tmax) and 607-639 (trange) */
    TB_JETS_DECL
    TB_COMPUTE_JETS(T); /* block-lockstep build (TB_BAR): idle lanes run the jets too, for the barriers inside */
#if TB_TMEM && TB_FINAL_ONLY
    /* tensor-memory kernels: state jets may live in TMEM, whose loads are warp-collective, so the step size and Horner
     * run on every lane of the (converged) warp; idle lanes and zero steps are sorted out afterwards                 */
    double dt = TB_STEPSIZE(a);
    const bool zero_step = dt == 0.0;
    const double rem = B_(T.tmax, T.t0);
    dt = M_(T.sign, TB_DTMIN(dt, M_(T.sign, rem)));
    __syncwarp();
    double xn[TB_DOF];
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) xn[c] = TB_HORNER(c, dt);
    if (!live) return false;
    if (zero_step) { T.st |= TB_ST_ZERO_STEP; return true; }
#else
    if (!live) return false;
    double dt = TB_STEPSIZE(a);
    if (dt == 0.0) { T.st |= TB_ST_ZERO_STEP; return true; }
    const double rem = B_(T.tmax, T.t0);
    dt = M_(T.sign, TB_DTMIN(dt, M_(T.sign, rem)));
    double xn[TB_DOF];
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) xn[c] = TB_HORNER(c, dt);
#endif
#if TB_FINAL_ONLY
    /* final-state build (TB_CF_FINAL_ONLY): no output grid, no per-step or dense output in the step body; the host
     * refuses such requests for this kernel.  Code that is only disabled at run time still costs the jets' scheduler
     * registers (measured on the Lorenz column kernel: 6 %).                                                     */
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) T.x0[c] = xn[c];
    T.t0 = A_(T.t0, dt);
    tb_update_cache(smt, T.x0);
    T.nsteps += 1;
#else
    if (a.trange) {
        const double tnext = A_(T.t0, dt);
        const long long N = a.N;
        while (M_(T.sign, T.t1) < M_(T.sign, tnext)) {
            const double d1 = B_(T.t1, T.t0);
            if (a.xv_trange) {
                TB_UNROLL_V2
                for (int c = 0; c < TB_DOF; c++)
                    a.xv_trange[((long long)c * a.nt + (T.iter - 1)) * N + T.i] = TB_HORNER(c, d1);
            }
            T.iter += 1;
            T.t1 = a.trange[T.iter - 1];
        }
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) T.x0[c] = xn[c];
        if (dt == rem) {
            if (a.xv_trange)
                for (int c = 0; c < TB_DOF; c++) a.xv_trange[((long long)c * a.nt + (T.iter - 1)) * N + T.i] = T.x0[c];
            T.t0 = tnext;
            return true;
        }
        T.t0 = tnext;
        tb_update_cache(smt, T.x0);
        T.nsteps += 1;
    } else {
        tb_store_psol(T, TB_JETS_ARGS, a);
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) T.x0[c] = xn[c];
        T.t0 = A_(T.t0, dt);
        tb_update_cache(smt, T.x0);
        T.nsteps += 1;
        tb_store_step(T, smt, a);
    }
#endif /* TB_FINAL_ONLY */
    if (T.nsteps > a.maxsteps) { T.st |= TB_ST_MAXSTEPS; return true; }
    if (!tb_isfinite_state(T.x0)) { T.st |= TB_ST_NONFINITE; return true; }
    return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
#elif TB_KIND == 1
    /* taylorinteg!(dense, f!, g, ...): src/rootfinding.jl:392-436; findroot! 176-235 */
    const int eo = a.eventorder;
    TB_JETS_DECL
    T.dt_old = T.dt;
    TB_COMPUTE_JETS(T);
    double dt = TB_STEPSIZE(a);
    dt = M_(T.sign, TB_DTMIN(dt, M_(T.sign, B_(T.tmax, T.t0))));
    T.dt = dt;
    double xn[TB_DOF];
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) xn[c] = TB_HORNER(c, dt);
#if !TB_FINAL_ONLY
    if (a.trange) {
        /* the trange variant with events, src/rootfinding.jl:530-549: in-step outputs first; the step that lands
         * on tmax stores its end point and leaves the loop BEFORE the event test of that step                  */
        const double tnext = A_(T.t0, dt);
        const long long N = a.N;
        while (M_(T.sign, T.t1) < M_(T.sign, tnext)) {
            const double d1 = B_(T.t1, T.t0);
            if (a.xv_trange) {
                TB_UNROLL_V2
                for (int c = 0; c < TB_DOF; c++)
                    a.xv_trange[((long long)c * a.nt + (T.iter - 1)) * N + T.i] = TB_HORNER(c, d1);
            }
            T.iter += 1;
            T.t1 = a.trange[T.iter - 1];
        }
        if (dt == B_(T.tmax, T.t0)) {
            if (a.xv_trange)
                for (int c = 0; c < TB_DOF; c++) a.xv_trange[((long long)c * a.nt + (T.iter - 1)) * N + T.i] = xn[c];
#pragma unroll
            for (int c = 0; c < TB_DOF; c++) T.x0[c] = xn[c];
            T.t0 = tnext;
            return true;
        }
    } else
        tb_store_psol(T, TB_JETS_ARGS, a);
#endif
#if TB_VARIANT == 1
    const bool f_now = tb_eval_event(smt, T.par, T.t0);
    const double g_now = TB_EV_G(eo);
#else
    const bool f_now = TB_EV_GATE_EXPR;
    const double g_now = tb_u_gcoef(J, smt, x0, eo);
#endif
    if (T.f_old && f_now && (M_(T.g_old, g_now) < 0.0)) { /* surfacecrossing: rootfinding.jl:124-132 */
#if TB_VARIANT == 2
        tb_u_gcopy(J, smt, x0); /* the series of g into its shared-memory slot: findroot! indexes it dynamically */
#endif
        int nriter = 1;
        const double slope = D_(B_(g_now, T.g_old), T.dt_old);
        const double dt_li = N_(D_(g_now, slope));
        double dt_nr = dt_li, gv, dgv;
        tb_eval_g_dg(smt, eo, dt_nr, &gv, &dgv);
        while (fabs(gv) > a.nrabstol && nriter <= a.newtoniter) { /* nrconvergencecriterion 147-152 */
            dt_nr = B_(dt_nr, D_(gv, dgv));
            tb_eval_g_dg(smt, eo, dt_nr, &gv, &dgv);
            nriter += 1;
        }
        if (nriter == a.newtoniter + 1) T.st |= TB_ST_NR_NOCONV;
        if (T.nevents <= a.event_cap) {
            const long long e = T.nevents - 1, N = a.N;
            if (a.tevents) a.tevents[e * N + T.i] = A_(T.t0, dt_nr);
            if (a.xevents) {
                TB_UNROLL_V2
                for (int c = 0; c < TB_DOF; c++)
                    a.xevents[((long long)c * a.event_cap + e) * N + T.i] = TB_HORNER(c, dt_nr);
            }
            if (a.gresids) a.gresids[e * N + T.i] = gv;
        } else
            T.st |= TB_ST_EVENT_OVERFLOW;
        T.nevents += 1;
    }
#pragma unroll
    for (int c = 0; c < TB_DOF; c++) T.x0[c] = xn[c];
    T.g_old = g_now;
    T.f_old = f_now;
    T.t0 = A_(T.t0, dt);
    tb_update_cache(smt, T.x0);
    T.nsteps += 1;
#if !TB_FINAL_ONLY
    tb_store_step(T, smt, a);
#endif
    if (T.nsteps > a.maxsteps) { T.st |= TB_ST_MAXSTEPS; return true; }
    if (!tb_isfinite_state(T.x0)) { T.st |= TB_ST_NONFINITE; return true; }
    return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
#else
    /* lyap_taylorinteg!: src/lyapunovspectrum.jl:379-426; lyap_taylorstep! 226-274 */
    tb_jetcoeffs(smt, T.par, T.t0);
#if TB_JAC_NOPS > 0
    tb_jac_k0(smt, T.par, T.t0);
#pragma unroll 1
    for (int k = 1; k < TB_P; k++) tb_jac_k(smt, T.par, T.t0, k);
#endif
    /* lyap_jetcoeffs!: lyapunovspectrum.jl:161-211 (generated, unrolled over i3,i2,i1) */
    tb_var_k0(smt, T.par);
#pragma unroll 1
    for (int k = 1; k < TB_P; k++) tb_var_k(smt, T.par, k);
    double dt = tb_stepsize(smt, a);
    dt = M_(T.sign, TB_DTMIN(dt, M_(T.sign, B_(T.tmax, T.t0))));
#pragma unroll
    for (int c = 0; c < TB_NX; c++) T.x0[c] = tb_horner(smt, c, dt);
    double QH[TB_DOF * TB_DOF], Rd[TB_DOF];
    tb_mgs(&T.x0[TB_DOF], QH, Rd);
    T.t0 = A_(T.t0, dt);
    const double tspan = B_(T.t0, T.t00);
    T.nsteps += 1;
    {
        const long long s = T.nsteps - 1, N = a.N;
        const bool st_ok = s < a.stride;
        if (a.tv && st_ok) a.tv[s * N + T.i] = T.t0;
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) {
            if (a.xv && st_ok) a.xv[((long long)c * a.stride + s) * N + T.i] = T.x0[c];
            T.lsum[c] = A_(T.lsum[c], tb_log(Rd[c]));
            T.lam[c] = D_(T.lsum[c], tspan);
            if (a.lambda_v && st_ok) a.lambda_v[((long long)c * a.stride + s) * N + T.i] = T.lam[c];
        }
    }
#pragma unroll
    for (int c = 0; c < TB_DOF * TB_DOF; c++) T.x0[TB_DOF + c] = QH[c];
    tb_update_cache(smt, T.x0);
    if (T.nsteps > a.maxsteps) { T.st |= TB_ST_MAXSTEPS; return true; }
    if (!tb_isfinite_state(T.x0)) { T.st |= TB_ST_NONFINITE; return true; }
    return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
#endif
}

__device__ __forceinline__ void tb_traj_finish(const TbTraj &T, const TbArgs &a) {
    const long long N = a.N, i = T.i;
    if (a.xfinal)
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) a.xfinal[(long long)c * N + i] = T.x0[c];
    if (a.tfinal) a.tfinal[i] = T.t0;
    if (a.nsteps) a.nsteps[i] = T.nsteps - 1;
    if (a.status) a.status[i] = T.st;
#if TB_KIND == 1
    if (a.nevents) a.nevents[i] = (T.nevents - 1 < a.event_cap) ? (T.nevents - 1) : a.event_cap;
#endif
#if TB_KIND == 2
    if (a.lambda)
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) a.lambda[(long long)c * N + i] = T.lam[c];
    if (a.lyap_state) {
#pragma unroll
        for (int c = 0; c < TB_DOF * TB_DOF; c++) a.lyap_state[(long long)c * N + i] = T.x0[TB_DOF + c];
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) a.lyap_state[(long long)(TB_DOF * TB_DOF + c) * N + i] = T.lsum[c];
        a.lyap_state[(long long)(TB_DOF * TB_DOF + TB_DOF) * N + i] = T.t00;
    }
#endif
}

/* ---- the fused ensemble kernel ---------------------------------------------
 * Persistent lanes: the grid is sized to the machine (one CTA of TB_BT threads
 * per SM slot), and every lane pulls trajectory indices from a global queue.
 * A lane that finishes refills immediately, so lanes of a warp always execute
 * the same step code whatever their step counts are (no tail of idle lanes);
 * the loop ends on a warp vote.  With a.lockstep a warp refills only when all
 * of its lanes are idle and takes 32 consecutive trajectories, which keeps the
 * per-step stores of tv/xv/psol coalesced.                                    */
extern "C" __global__ void __launch_bounds__(TB_BT, TB_MINB) tb_ensemble_kernel(const TbArgs a) {
    extern __shared__ double tb_sm[];
    double *__restrict__ smt = tb_sm + threadIdx.x;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
#if TB_TMEM
    __shared__ unsigned tb_tm_slot;
    const unsigned tb_tm = tb_tm_alloc(&tb_tm_slot);
#else
    const unsigned tb_tm = 0u;
#endif
    TbTraj T;
#if TB_BAR || TB_TMEM
    /* idle lanes execute the jet code on whatever T holds: give it defined numbers */
#pragma unroll
    for (int c = 0; c < TB_NX; c++) T.x0[c] = 1.0;
#pragma unroll
    for (int j = 0; j < (TB_NPAR > 0 ? TB_NPAR : 1); j++) T.par[j] = 1.0;
    T.t0 = 0.0;
#endif
    bool have = false, more = true;
    for (;;) {
        if (a.lockstep) {
            if (more && !__any_sync(full, have)) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(a.counter, 32ull);
                base = __shfl_sync(full, base, 0);
                if (base >= (unsigned long long)a.N) more = false;
                else {
                    long long i = (long long)base + lane;
                    if (i < a.N) {
                        have = true;
                        if (a.queue_reverse) i = a.N - 1 - i;
                        if (tb_traj_init(T, smt, a, i)) { tb_traj_finish(T, a); have = false; }
                    }
                }
            }
        } else {
            while (!have && more) {
                long long i = (long long)atomicAdd(a.counter, 1ull);
                if (i >= a.N) { more = false; break; }
                have = true;
                /* a lane that refills near the end of the queue runs alone while the others idle: when the caller
                 * says the cost grows with the index, the expensive trajectories are handed out first            */
                if (a.queue_reverse) i = a.N - 1 - i;
                if (tb_traj_init(T, smt, a, i)) { tb_traj_finish(T, a); have = false; }
            }
        }
#if TB_BAR
        /* block lockstep: every warp runs every step until the whole block is out of work, so the
         * barriers inside the jet code are uniform and the warps share one instruction stream  */
        if (!__syncthreads_or(have || (a.lockstep && more))) break;
#if TB_BAR == 2
        /* step lockstep: one barrier per step only; the warps start every step together and drift by
         * less than the instruction cache holds, so a line fetched by the leader serves the others */
        if (have) { if (tb_traj_step(T, smt, tb_tm, a, true)) { tb_traj_finish(T, a); have = false; } }
#else
        if (tb_traj_step(T, smt, tb_tm, a, have)) { tb_traj_finish(T, a); have = false; }
#endif
#else
        if (!__any_sync(full, have)) { if (!(a.lockstep && more)) break; else continue; }
#if TB_TMEM
        /* the jet body holds warp-collective tensor-memory accesses: every lane of the warp runs it (the vote above
         * reconverged the warp), idle lanes on stale numbers whose results are dropped                             */
        if (tb_traj_step(T, smt, tb_tm, a, have)) { tb_traj_finish(T, a); have = false; }
#else
        if (have) {
            if (tb_traj_step(T, smt, tb_tm, a, true)) { tb_traj_finish(T, a); have = false; }
        }
#endif
#endif
    }
#if TB_TMEM
    tb_tm_free(&tb_tm_slot);
#endif
}

#else /* TB_KIND == 3 */
/* ---- column-parallel Lyapunov spectrum (TB_KIND == 3) ---------------------------------------------
 * lyap_taylorinteg!: src/lyapunovspectrum.jl:340-429; lyap_taylorstep! 226-274; lyap_jetcoeffs! 161-211.
 * A TEAM of TB_NQ consecutive lanes integrates one trajectory.  Lane j carries the state x (all lanes
 * compute identical x jets) and column j of the fundamental matrix; the generated jets are those of the
 * column system [x; v], v' = J(x) v, whose sums follow the reference's accumulation order, so every
 * coefficient has the bits of the single-thread kernel (TB_KIND == 2) and of the oracle.  Per step the
 * columns are exchanged with warp shuffles and every lane runs the same modified Gram-Schmidt.       */
#define TB_TEAM TB_NQ
#define TB_TPW (32 / TB_TEAM) /* teams per warp */

__device__ __forceinline__ bool tb_traj_init(TbTraj &T, double *__restrict__ smt, const TbArgs &a, long long i) {
    T.i = i;
    const long long N = a.N;
#pragma unroll
    for (int c = 0; c < TB_NQ; c++) T.x0[c] = a.q0[(long long)c * N + i];
    /* init_cache_lyap: state = [q; vec(I)]  src/integrator/cache.jl:320-369; this lane holds column `lit` */
#pragma unroll
    for (int r = 0; r < TB_NQ; r++) T.x0[TB_NQ + r] = (r == T.lit) ? 1.0 : 0.0;
    T.lsum = 0.0;
    T.lam = 0.0;
    if (a.lyap_state0) { /* continue from state: this lane's column of Q and its running sum (taylorb200.h) */
#pragma unroll
        for (int r = 0; r < TB_NQ; r++) T.x0[TB_NQ + r] = a.lyap_state0[(long long)(T.lit * TB_NQ + r) * N + i];
        T.lsum = a.lyap_state0[(long long)(TB_NQ * TB_NQ + T.lit) * N + i];
    }
#pragma unroll
    for (int j = 0; j < TB_NPAR; j++) T.par[j] = a.params_per_traj ? a.params[(long long)j * N + i] : a.params[j];
    T.nsteps = 1;
    T.st = 0;
    T.iter = 2;
#if !TB_LYAP_TRANGE
    if (false) {
#else
    if (a.trange) { /* src/lyapunovspectrum.jl:504-515; xv and lambda NaN-prefilled (cache.jl:421-422) */
#endif
        T.t0 = a.trange[0];
        T.t1 = a.trange[1];
        T.tmax = a.trange[a.nt - 1];
        if (a.xv_trange && T.lit == 0)
            for (int c = 0; c < TB_NQ; c++) {
                a.xv_trange[((long long)c * a.nt) * N + i] = T.x0[c];
                for (long long j = 1; j < a.nt; j++) a.xv_trange[((long long)c * a.nt + j) * N + i] = TB_NAN;
            }
        if (a.lambda_v) {
            a.lambda_v[((long long)T.lit * a.nt) * N + i] = 0.0;
            for (long long j = 1; j < a.nt; j++) a.lambda_v[((long long)T.lit * a.nt + j) * N + i] = TB_NAN;
        }
    } else {
        T.t0 = a.t0;
        T.tmax = a.tmax;
        T.t1 = 0.0;
    }
    T.sign = copysign(1.0, B_(T.tmax, T.t0));
    T.t00 = a.lyap_state0 ? a.lyap_state0[(long long)(TB_NQ * TB_NQ + TB_NQ) * N + i] : T.t0;
#if TB_LYAP_TRANGE
    if (a.trange) return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
#endif
    if (T.lit == 0) {
        if (a.tv) a.tv[i] = T.t0;
        if (a.xv)
#pragma unroll
            for (int c = 0; c < TB_NQ; c++) a.xv[((long long)c * a.stride) * N + i] = T.x0[c];
    }
    if (a.lambda_v) a.lambda_v[((long long)T.lit * a.stride) * N + i] = 0.0;
    return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
}

/* value held by column `col` of a per-lane array indexed [r + col * TB_NQ], col known only at run time */
__device__ __forceinline__ double tb_pick_col(const double *M, int r, int col) {
    double v = M[r];
#pragma unroll
    for (int j = 1; j < TB_NQ; j++) v = (col == j) ? M[r + j * TB_NQ] : v;
    return v;
}

__device__ __forceinline__ bool tb_traj_step(TbTraj &T, double *__restrict__ smt, const unsigned tb_tm, const TbArgs &a, const bool live) {
    TB_JETS_DECL
    TB_COMPUTE_JETS(T);
    double dt = TB_STEPSIZE(a); /* over the TB_NQ components of x only (lyapunovspectrum.jl:262) */
    dt = M_(T.sign, TB_DTMIN(dt, M_(T.sign, B_(T.tmax, T.t0))));
    double xn[TB_NX];
    TB_UNROLL_V2
    for (int c = 0; c < TB_NX; c++) xn[c] = TB_HORNER(c, dt);
    /* the evolved fundamental matrix, column j from lane leader + j */
    double A[TB_NQ * TB_NQ], QH[TB_NQ * TB_NQ], Rd[TB_NQ];
#pragma unroll
    for (int j = 0; j < TB_NQ; j++)
#pragma unroll
        for (int r = 0; r < TB_NQ; r++) A[r + j * TB_NQ] = __shfl_sync(T.tmask, xn[TB_NQ + r], T.leader + j);
    tb_mgs(A, QH, Rd);
    double rd = Rd[0];
#pragma unroll
    for (int j = 1; j < TB_NQ; j++) rd = (T.lit == j) ? Rd[j] : rd;
    const double lr = tb_log(rd);
#if TB_LYAP_TRANGE /* compiled in only for kernels built with TB_CF_LYAP_TRANGE: the extra live values cost the
                    * (t0, tmax) driver 6 % on Lorenz-28 */
    if (a.trange) {
        /* in-step outputs of the trange variant, src/lyapunovspectrum.jl:545-571: states at t1, exponents from the
         * end-of-step fundamental matrix, the reference's literal `t1 < tnext` test; the step that lands on tmax
         * stores its end point and leaves the loop                                                             */
        const double tnext = A_(T.t0, dt);
        const long long N = a.N;
        while (T.t1 < tnext) {
            const double d1 = B_(T.t1, T.t0);
            if (a.xv_trange && T.lit == 0) {
                TB_UNROLL_V2
                for (int c = 0; c < TB_NQ; c++)
                    a.xv_trange[((long long)c * a.nt + (T.iter - 1)) * N + T.i] = TB_HORNER(c, d1);
            }
            if (a.lambda_v) a.lambda_v[((long long)T.lit * a.nt + (T.iter - 1)) * N + T.i] = D_(A_(T.lsum, lr), B_(T.t1, T.t00));
            T.iter += 1;
            T.t1 = a.trange[T.iter - 1];
        }
        if (dt == B_(T.tmax, T.t0)) {
            T.lam = D_(A_(T.lsum, lr), B_(T.tmax, T.t00));
            if (a.xv_trange && T.lit == 0)
                for (int c = 0; c < TB_NQ; c++) a.xv_trange[((long long)c * a.nt + (T.iter - 1)) * N + T.i] = xn[c];
            if (a.lambda_v) a.lambda_v[((long long)T.lit * a.nt + (T.iter - 1)) * N + T.i] = T.lam;
#pragma unroll
            for (int c = 0; c < TB_NQ; c++) T.x0[c] = xn[c];
            T.t0 = tnext;
            return true;
        }
    }
#endif
    T.t0 = A_(T.t0, dt);
    const double tspan = B_(T.t0, T.t00);
    T.nsteps += 1;
    T.lsum = A_(T.lsum, lr);
    T.lam = D_(T.lsum, tspan);
    {
        const long long s = T.nsteps - 1, N = a.N;
        const bool st_ok = s < a.stride;
        if (T.lit == 0 && st_ok) {
            if (a.tv) a.tv[s * N + T.i] = T.t0;
            if (a.xv)
#pragma unroll
                for (int c = 0; c < TB_NQ; c++) a.xv[((long long)c * a.stride + s) * N + T.i] = xn[c];
        }
#if TB_LYAP_TRANGE
        const bool per_step = !a.trange; /* in the trange variant lambda_v has the [dof x nt x N] layout */
#else
        const bool per_step = true;
#endif
        if (a.lambda_v && st_ok && per_step) a.lambda_v[((long long)T.lit * a.stride + s) * N + T.i] = T.lam;
    }
#pragma unroll
    for (int c = 0; c < TB_NQ; c++) T.x0[c] = xn[c];
#pragma unroll
    for (int r = 0; r < TB_NQ; r++) T.x0[TB_NQ + r] = tb_pick_col(QH, r, T.lit);
    if (T.nsteps > a.maxsteps) { T.st |= TB_ST_MAXSTEPS; return true; }
    if (!tb_isfinite_state(T.x0)) { T.st |= TB_ST_NONFINITE; return true; }
    return !(M_(T.sign, T.t0) < M_(T.sign, T.tmax));
}

__device__ __forceinline__ void tb_traj_finish(const TbTraj &T, const TbArgs &a) {
    const long long N = a.N, i = T.i;
    if (T.lit == 0) {
        if (a.xfinal)
#pragma unroll
            for (int c = 0; c < TB_NQ; c++) a.xfinal[(long long)c * N + i] = T.x0[c];
        if (a.tfinal) a.tfinal[i] = T.t0;
        if (a.nsteps) a.nsteps[i] = T.nsteps - 1;
        if (a.status) a.status[i] = T.st;
    }
    if (a.lambda) a.lambda[(long long)T.lit * N + i] = T.lam;
    if (a.lyap_state) {
#pragma unroll
        for (int r = 0; r < TB_NQ; r++) a.lyap_state[(long long)(T.lit * TB_NQ + r) * N + i] = T.x0[TB_NQ + r];
        a.lyap_state[(long long)(TB_NQ * TB_NQ + T.lit) * N + i] = T.lsum;
        if (T.lit == 0) a.lyap_state[(long long)(TB_NQ * TB_NQ + TB_NQ) * N + i] = T.t00;
    }
}

/* Persistent teams: every team pulls trajectory indices from the global queue (leader lane does the
 * atomic, the index is broadcast inside the team) and refills as soon as its trajectory ends; lanes
 * beyond the last whole team of a warp (32 % TB_NQ of them) stay idle.  The loop ends on a warp vote. */
extern "C" __global__ void __launch_bounds__(TB_BT, TB_MINB) tb_ensemble_kernel(const TbArgs a) {
    extern __shared__ double tb_sm[];
    double *__restrict__ smt = tb_sm + threadIdx.x;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool member = lane < TB_TPW * TB_TEAM;
    const unsigned tb_tm = 0u; /* no tensor-memory tier in the team kernel */
    TbTraj T;
    T.lit = lane % TB_TEAM;
    T.leader = lane - T.lit;
    T.tmask = member ? ((TB_TEAM == 32 ? 0xffffffffu : ((1u << TB_TEAM) - 1u)) << T.leader) : (1u << lane);
    bool have = false, more = member;
    for (;;) {
        while (!have && more) {
            long long i = 0;
            if (T.lit == 0) i = (long long)atomicAdd(a.counter, 1ull);
            i = __shfl_sync(T.tmask, i, T.leader);
            if (i >= a.N) { more = false; break; }
            have = true;
            if (a.queue_reverse) i = a.N - 1 - i;
            if (tb_traj_init(T, smt, a, i)) { tb_traj_finish(T, a); have = false; }
        }
#if TB_BAR == 2
        if (!__syncthreads_or(have)) break; /* step lockstep: the warps of the block share the instruction stream */
#else
        if (!__any_sync(full, have)) break;
#endif
        if (have) {
            if (tb_traj_step(T, smt, tb_tm, a, true)) { tb_traj_finish(T, a); have = false; }
        }
    }
}
#endif /* TB_KIND == 3 */

#if TB_KIND != 3
/* tb_jet: one taylorstep! (src/integrator/taylorstep.jl:59-82) per trajectory;
 * jets x[dof x (P+1) x N] and the step size.                                   */
extern "C" __global__ void __launch_bounds__(TB_BT, TB_MINB) tb_jet_kernel(const TbArgs a) {
    extern __shared__ double tb_sm[];
    double *__restrict__ smt = tb_sm + threadIdx.x;
#if TB_TMEM
    __shared__ unsigned tb_tm_slot;
    const unsigned tb_tm = tb_tm_alloc(&tb_tm_slot);
#else
    const unsigned tb_tm = 0u;
#endif
    /* whole warps iterate (the jet body may hold warp-collective accesses); lanes past N recompute trajectory N-1 */
    for (long long i0 = (long long)blockIdx.x * TB_BT + (threadIdx.x & ~31); i0 < a.N; i0 += (long long)gridDim.x * TB_BT) {
        const bool live = i0 + (threadIdx.x & 31) < a.N;
        const long long i = live ? i0 + (threadIdx.x & 31) : a.N - 1;
        double par[TB_NPAR > 0 ? TB_NPAR : 1];
        for (int j = 0; j < TB_NPAR; j++) par[j] = a.params_per_traj ? a.params[(long long)j * a.N + i] : a.params[j];
        struct { double x0[TB_DOF]; double par[TB_NPAR > 0 ? TB_NPAR : 1]; double t0; } T;
        for (int j = 0; j < TB_NPAR; j++) T.par[j] = par[j];
        T.t0 = a.t0;
#pragma unroll
        for (int c = 0; c < TB_DOF; c++) T.x0[c] = a.q0[(long long)c * a.N + i];
        tb_update_cache(smt, T.x0);
        TB_JETS_DECL
        TB_COMPUTE_JETS(T);
        double dt = TB_STEPSIZE(a);
        if (a.dtout && live) a.dtout[i] = dt;
        if (a.xjet) {
            TB_UNROLL_V2
            for (int c = 0; c < TB_DOF; c++) {
                TB_UNROLL_V2
                for (int k = 0; k <= TB_P; k++) {
                    const double v = TB_XCOEF(c, k); /* read by every lane (may be a warp-collective load) */
                    if (live) a.xjet[((long long)c * TB_NP1 + k) * a.N + i] = v;
                }
            }
        }
    }
#if TB_TMEM
    tb_tm_free(&tb_tm_slot);
#endif
}

#else
/* the jet probe is not offered for the column system (tb_jet returns TB_ERR_UNSUPPORTED) */
extern "C" __global__ void tb_jet_kernel(const TbArgs a) {}
#endif

#endif /* TB_DEVICE_DRV_CUH */
