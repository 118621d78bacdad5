/* tb_wide2_dev.cuh -- device side of the cluster kernel for large-state ODEs (see tb_wide2.cu for the design).
 * Compiled twice: by nvcc into libtaylorb200.so as the INTERPRETER (entries read from shared/global memory), and by
 * NVRTC per program with TB_W2_GENERATED defined: the group programs are emitted as straight-line calls of exec2 with
 * compile-time entries (tb_wide2.cu, wide2_generate), so the dispatch, the operand arithmetic and the flag tests fold
 * away and the coupling sum is unrolled with its length known.  No #include here: the text is embedded for NVRTC. */
#ifndef TB_WIDE2_DEV_CUH
#define TB_WIDE2_DEV_CUH

/* location of a series: >= 0 persistent id (shared memory W[k*NS + id]); -1 a number; <= -2 transient slot -2 - loc */
struct C2Op {
    int op, res, a, b, aux;
    int flags; /* bit 0: operand a enters negated, bit 1: operand b */
    int para, parb;
    int lc_off, lc_n, lc_stride, lc_soff, lc_sstride;
    int pad;
    double imma, immb, imm;
};

struct W2Args {
    long long N;
    int dof, npar, P, NS, ncl, maxlen, ntr, tsid;
    const C2Op *ops; /* [maxlen x ncl], op-major */
    const double *lc_coef;
    const int *lc_ser;
    const int *xsid; /* [dof] persistent id of x[c] */
    const int *cl_len; /* [ncl] entries of cluster c (0: a padding slot) */
    int lc_smem, lc_count; /* coupling constants copied to shared memory (lc_count doubles) */
    int ngroups;           /* > 0: group programs */
    int g_smem, ngops, ngnum; /* group programs copied to shared memory */
    int g_first[8], g_count[8], g_off[8], g_len[8];
    const C2Op *gops;
    const double *gnum;
    const double *q0, *params, *trange;
    long long nt;
    double t0, tmax, abstol, reltol, minstep, maxstep;
    long long maxsteps;
    int params_per_traj;
    double *xfinal, *tfinal;
    long long *nsteps;
    int *status;
    double *xv_trange;
    unsigned long long *counter;
};

#ifndef TB_W2_LCB
#define TB_W2_LCB 8 /* terms per batch of the coupling sum */
#endif
#ifdef TB_W2_GENERATED
#define TB_W2_LC_UNROLL _Pragma("unroll")
#else
#define TB_W2_LC_UNROLL _Pragma("unroll 1")
#endif
#define XM(a, b) __dmul_rn((a), (b))
#define XA(a, b) __dadd_rn((a), (b))
#define XS(a, b) __dsub_rn((a), (b))
#define XD(a, b) __ddiv_rn((a), (b))
template <bool FAST> __device__ __forceinline__ double mad2(double a, double b, double c) {
    return FAST ? fma(a, b, c) : __dadd_rn(c, __dmul_rn(a, b));
}

/* x / d for the small integers of solcoeff! (TaylorIntegration.jl:33-40) as three FP64 instructions: y = RN(1/d) from a
 * table filled once per CTA, q = RN(x y), r = x - q d (exact in one FMA), RN(q + r y) = RN(x / d) (Markstein) when
 * nothing under/overflows; operands outside [2^-767, 2^768) (and 0, Inf, NaN) take the IEEE division.  Same bits. */
__shared__ double w2_rinv[64];
/* out of line on purpose: inlined, ptxas if-converts the rare branch and every call pays the whole division */
__device__ __noinline__ double w2_div_ieee(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double w2_div_int(double x, int d) {
    if (d >= 64) return w2_div_ieee(x, (double)d);
    const double y = w2_rinv[d], dd = (double)d;
    const double q = XM(x, y);
    const double r = __fma_rn(-q, dd, x);
    double q2 = __fma_rn(r, y, q);
    const unsigned e = ((unsigned)__double2hiint(x) & 0x7ff00000u) - 0x10000000u;
    if (!(e < 0x5ff00000u)) q2 = w2_div_ieee(x, dd);
    return q2;
}

/* one entry of a cluster at order k: the recurrences of SURVEY App. B in the oracle's order (as tb_wide.cu exec_op) */
template <bool FAST>
__device__ __forceinline__ void exec2(const W2Args &A, double *__restrict__ W, double *__restrict__ tr, const double *par,
                                      const double *__restrict__ lcc, const bool lc_smem, const C2Op &o, const int k,
                                      const double *creg = nullptr, const int ncreg = 0) {
#ifdef TB_W2_GENERATED
    const int NS = TB_W2_NS, TRS = TB_W2_THREADS;
#else
    const int NS = A.NS, TRS = blockDim.x;
#endif
    const bool sa = o.a == -1, sb = o.b == -1;
    const double na = o.para >= 0 ? par[o.para] : o.imma, nb = o.parb >= 0 ? par[o.parb] : o.immb;
    /* value of operand series `loc` at index j (transients exist at j == k only) */
    auto rd = [&](int loc, int j) -> double { return loc >= 0 ? W[j * NS + loc] : tr[(-2 - loc) * TRS]; };
    int op = o.op;
    if (op == TB_OP_POW) {
        if (o.imm == 0.0) op = TB_OP_ONE;
        else if (o.imm == 1.0) op = TB_OP_IDENTITY;
        else if (o.imm == 2.0) op = TB_OP_SQR;
        else if (o.imm == 0.5) op = TB_OP_SQRT;
    }
    double r = 0.0;
    switch (op) {
        case W_NOP: return;
        case W_SOLCOEFF: /* x[k+1] = dx[k] / (k+1): TaylorIntegration.jl:33-40 */
            W[(k + 1) * NS + o.res] = FAST ? rd(o.a, k) * (1.0 / (double)(k + 1)) : w2_div_int(rd(o.a, k), k + 1);
            return;
        case TB_OP_IDENTITY: r = sa ? (k == 0 ? na : 0.0) : rd(o.a, k); break;
        case TB_OP_NEG: r = -rd(o.a, k); break;
        case TB_OP_ZERO: r = 0.0; break;
        case TB_OP_ONE: r = k == 0 ? 1.0 : 0.0; break;
        case TB_OP_ADD:
            if (sa && sb) r = k == 0 ? XA(na, nb) : 0.0;
            else if (sa) r = k == 0 ? XA(na, rd(o.b, 0)) : rd(o.b, k);
            else if (sb) r = k == 0 ? XA(rd(o.a, 0), nb) : rd(o.a, k);
            else r = XA(rd(o.a, k), rd(o.b, k));
            break;
        case TB_OP_SUB:
            if (sa && sb) r = k == 0 ? XS(na, nb) : 0.0;
            else if (sa) r = k == 0 ? XS(na, rd(o.b, 0)) : -rd(o.b, k);
            else if (sb) r = k == 0 ? XS(rd(o.a, 0), nb) : rd(o.a, k);
            else r = XS(rd(o.a, k), rd(o.b, k));
            break;
        case TB_OP_MUL:
            if (sa && sb) r = k == 0 ? XM(na, nb) : 0.0;
            else if (sa) r = XM(na, rd(o.b, k));
            else if (sb) r = XM(rd(o.a, k), nb);
            else { /* Cauchy product; a negated operand is negated on the fly (exact: a sign bit on the DMUL operand) */
                const double *pa = W + o.a, *pb = W + o.b + k * NS;
                const bool neg = ((o.flags & 1) != 0) != ((o.flags & 2) != 0);
                r = XM(neg ? -pa[0] : pa[0], pb[0]);
                if (FAST) {
                    for (int i = 1; i <= k; i++) { pa += NS; pb -= NS; r = fma(neg ? -pa[0] : pa[0], pb[0], r); }
                } else {
                    int i = 1;
                    for (; i + 2 <= k + 1; i += 2) { /* two products ahead of the dependent adds */
                        const double q0 = XM(pa[NS], pb[-NS]), q1 = XM(pa[2 * NS], pb[-2 * NS]);
                        pa += 2 * NS; pb -= 2 * NS;
                        r = XA(r, neg ? -q0 : q0); r = XA(r, neg ? -q1 : q1);
                    }
                    for (; i <= k; i++) { pa += NS; pb -= NS; const double q = XM(pa[0], pb[0]); r = XA(r, neg ? -q : q); }
                }
            }
            break;
        case W_MUL2: {
            /* two Cauchy products with the SAME second operand, fused by wide2_generate (many-spin: z_i * S_i and
             * -y_i * S_i): S is loaded once and the two dependent add chains interleave.  Each product keeps the
             * accumulation order of TB_OP_MUL above (same bits).  a2 in o.aux, res2 in o.lc_soff, bit 2: negate a2. */
            const double *pa = W + o.a, *pc = W + o.aux, *pb = W + o.b + k * NS;
            const bool bn = (o.flags & 2) != 0;
            const bool n1 = ((o.flags & 1) != 0) != bn, n2 = ((o.flags & 4) != 0) != bn;
            const double b0 = pb[0];
            double r1 = XM(n1 ? -pa[0] : pa[0], b0), r2 = XM(n2 ? -pc[0] : pc[0], b0);
            for (int i = 1; i <= k; i++) {
                pa += NS; pc += NS; pb -= NS;
                const double bb = pb[0];
                if (FAST) { r1 = fma(n1 ? -pa[0] : pa[0], bb, r1); r2 = fma(n2 ? -pc[0] : pc[0], bb, r2); }
                else {
                    const double q1 = XM(pa[0], bb), q2 = XM(pc[0], bb);
                    r1 = XA(r1, n1 ? -q1 : q1); r2 = XA(r2, n2 ? -q2 : q2);
                }
            }
            if (o.res >= 0) W[k * NS + o.res] = r1; else tr[(-2 - o.res) * TRS] = r1;
            if (o.lc_soff >= 0) W[k * NS + o.lc_soff] = r2; else tr[(-2 - o.lc_soff) * TRS] = r2;
            return;
        }
        case TB_OP_SQR: {
            const double *a = W + o.a;
            if (k == 0) r = XM(a[0], a[0]);
            else {
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                r = XM(a[0], a[k * NS]);
                for (int i = 1; i <= kend; i++) r = mad2<FAST>(a[i * NS], a[(k - i) * NS], r);
                r = XM(2.0, r);
                if (!kodd) r = mad2<FAST>(a[(k / 2) * NS], a[(k / 2) * NS], r);
            }
            break;
        }
        case TB_OP_DIV:
            if (sa && sb) r = k == 0 ? XD(na, nb) : 0.0;
            else if (sb) r = XD(rd(o.a, k), nb);
            else if (k == 0) r = XD(sa ? na : rd(o.a, 0), W[o.b]);
            else {
                const double *b = W + o.b, *c = W + o.res;
                double s = XM(c[0], b[k * NS]);
                for (int i = 1; i <= k - 1; i++) s = mad2<FAST>(c[i * NS], b[(k - i) * NS], s);
                r = sa ? XD(-s, b[0]) : XD(XS(rd(o.a, k), s), b[0]);
            }
            break;
        case TB_OP_SQRT: {
            const double *c = W + o.res;
            if (k == 0) r = __dsqrt_rn(rd(o.a, 0));
            else {
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                double s = 0.0;
                for (int i = 1; i <= kend; i++) s = i == 1 ? XM(c[i * NS], c[(k - i) * NS]) : mad2<FAST>(c[i * NS], c[(k - i) * NS], s);
                double aux = XS(rd(o.a, k), XM(2.0, s));
                if (!kodd) aux = XS(aux, XM(c[(kend + 1) * NS], c[(kend + 1) * NS]));
                r = XD(aux, XM(2.0, c[0]));
            }
            break;
        }
        case TB_OP_POW: {
            const double *a = W + o.a, *c = W + o.res;
            if (k == 0) r = tb_pow(a[0], o.imm);
            else {
                double s = XM(XM(XM(o.imm, (double)k), c[0]), a[k * NS]);
                for (int i = 1; i <= k - 1; i++) {
                    const double rr = XS(XM(o.imm, (double)(k - i)), (double)i);
                    s = mad2<FAST>(XM(rr, c[i * NS]), a[(k - i) * NS], s);
                }
                r = XD(s, XM((double)k, a[0]));
            }
            break;
        }
        case TB_OP_EXP: {
            const double *a = W + o.a, *c = W + o.res;
            if (k == 0) r = tb_exp(a[0]);
            else {
                double s = XM(XM((double)k, a[k * NS]), c[0]);
                for (int i = 1; i <= k - 1; i++) s = mad2<FAST>(XM((double)(k - i), a[(k - i) * NS]), c[i * NS], s);
                r = XD(s, (double)k);
            }
            break;
        }
        case TB_OP_LOG: {
            const double *a = W + o.a, *c = W + o.res;
            if (k == 0) r = tb_log(a[0]);
            else if (k == 1) r = XD(a[NS], a[0]);
            else {
                double s = XM(XM((double)(k - 1), a[NS]), c[(k - 1) * NS]);
                for (int i = 2; i <= k - 1; i++) s = mad2<FAST>(XM((double)(k - i), a[i * NS]), c[(k - i) * NS], s);
                r = XD(XS(a[k * NS], XD(s, (double)k)), a[0]);
            }
            break;
        }
        case TB_OP_SINCOS: {
            const double *a = W + o.a;
            double *sn = W + o.res, *cs = W + o.aux;
            if (k == 0) { double s, c; tb_sincos(a[0], &s, &c); sn[0] = s; cs[0] = c; return; }
            double x = a[NS];
            double sk = XM(x, cs[(k - 1) * NS]), ck = XM(-x, sn[(k - 1) * NS]);
            for (int i = 2; i <= k; i++) {
                x = XM((double)i, a[i * NS]);
                sk = mad2<FAST>(x, cs[(k - i) * NS], sk);
                ck = FAST ? fma(-x, sn[(k - i) * NS], ck) : XS(ck, XM(x, sn[(k - i) * NS]));
            }
            sn[k * NS] = XD(sk, (double)k);
            cs[k * NS] = XD(ck, (double)k);
            return;
        }
        case W_LINCOMB: { /* ((start + c1 x1[k]) + c2 x2[k]) + ... : the fused add!/mul! chain, in tape order */
            const double *cf = lcc + o.lc_off;
            const double *Wk = W + k * NS;
            const int ls = o.lc_stride, n = o.lc_n;
            int j = 0;
            if (o.lc_sstride == 0) {
                /* the terms are consecutive stored series (x_1[k], x_2[k], ...: the state jets of the other clusters):
                 * no id table, the operand pointer just advances; products are formed one batch ahead of the adds */
                const double *xs = Wk + o.lc_soff;
                const int xstep = o.pad;
                /* coefficient of term t: the first `ncreg` live in registers for the whole launch (generated kernels: the
                 * thread <-> cluster mapping is fixed and the constants do not change), the rest come from L1/L2 or smem */
#ifdef TB_W2_GENERATED
#define TB_W2_CF(t) ((t) < ncreg ? creg[(t)] : cf[(t) * ls])
#else
#define TB_W2_CF(t) cf[(t) * ls]
#endif
                if (sa) { const double m0 = XM(TB_W2_CF(0), xs[0]); r = k == 0 ? XA(na, m0) : m0; j = 1; }
                else r = rd(o.a, k);
                /* two stages one batch apart: the loads of batch b+1 are issued before the products and the dependent
                 * adds of batch b, so the chain never waits on shared memory / L2 (in-order issue) */
                constexpr int B = TB_W2_LCB;
                const int nb = (n - j) / B;
                if (nb > 0) {
                    double c[B], x[B];
#pragma unroll
                    for (int u = 0; u < B; u++) { c[u] = TB_W2_CF(j + u); x[u] = xs[(j + u) * xstep]; }
                    TB_W2_LC_UNROLL
                    for (int b = 0; b < nb; b++) {
                        double cn[B], xn[B];
                        j += B;
                        if (b + 1 < nb) {
#pragma unroll
                            for (int u = 0; u < B; u++) { cn[u] = TB_W2_CF(j + u); xn[u] = xs[(j + u) * xstep]; }
                        }
                        if (FAST) {
#pragma unroll
                            for (int u = 0; u < B; u++) r = fma(c[u], x[u], r);
                        } else {
                            double q[B];
#pragma unroll
                            for (int u = 0; u < B; u++) q[u] = XM(c[u], x[u]);
#pragma unroll
                            for (int u = 0; u < B; u++) r = XA(r, q[u]);
                        }
                        if (b + 1 < nb) {
#pragma unroll
                            for (int u = 0; u < B; u++) { c[u] = cn[u]; x[u] = xn[u]; }
                        }
                    }
                }
                TB_W2_LC_UNROLL
                for (; j < n; j++) r = mad2<FAST>(TB_W2_CF(j), xs[j * xstep], r);
#undef TB_W2_CF
            } else {
                const int *sr = A.lc_ser + o.lc_soff;
                const int ss = o.lc_sstride;
                if (sa) { const double m0 = XM(cf[0], Wk[__ldg(sr)]); r = k == 0 ? XA(na, m0) : m0; j = 1; cf += ls; sr += ss; }
                else r = rd(o.a, k);
                for (; j < n; j++, cf += ls, sr += ss) r = mad2<FAST>(cf[0], Wk[__ldg(sr)], r);
            }
            break;
        }
        default: break;
    }
    if (o.res >= 0) W[k * NS + o.res] = r;
    else tr[(-2 - o.res) * TRS] = r;
}

__device__ __forceinline__ double w2_horner(const double *X, int NS, int P, double dt) {
    double s = X[P * NS];
    for (int k = P - 1; k >= 0; k--) s = XA(XM(s, dt), X[k * NS]);
    return s;
}

/* block-wide min (or max) that PROPAGATES NaN like Julia's min / max (stepsize.jl:74, 86) */
__device__ double w2_block_reduce(double v, double *red, bool take_max) {
    bool nan = v != v;
    for (int o = 16; o > 0; o >>= 1) {
        const double u = __shfl_xor_sync(0xffffffffu, v, o);
        v = take_max ? fmax(v, u) : fmin(v, u);
    }
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    nan = __syncthreads_or(nan);
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < nw; i++) r = take_max ? fmax(r, red[i]) : fmin(r, red[i]);
    return nan ? __longlong_as_double(0x7FF8000000000000LL) : r;
}

#ifdef TB_W2_GENERATED
/* the group programs as straight-line code (emitted by wide2_generate after this header) */
template <bool FAST>
__device__ __forceinline__ void tb_w2_groups(const W2Args &a, double *__restrict__ W, double *__restrict__ tr, const double *par,
                                             const double *__restrict__ lcc, const double *__restrict__ gnum, const int k, const int tid,
                                             const double *lcreg);
/* coupling constants this thread keeps in registers for the whole launch (filled once; generated) */
__device__ __forceinline__ void tb_w2_lcreg_fill(double *lcreg, const double *__restrict__ lcc, const int tid);
#endif

template <bool FAST> __device__ __forceinline__ void wide2_body(const W2Args &a) {
    extern __shared__ double sm2[];
    __shared__ double red[8];
    __shared__ unsigned long long s_idx;
#ifdef TB_W2_GENERATED
    const int tid = threadIdx.x, nth = TB_W2_THREADS, P = TB_W2_P, NS = TB_W2_NS;
#else
    const int tid = threadIdx.x, nth = blockDim.x, P = a.P, NS = a.NS;
#endif
    double *W = sm2;                              /* [P+1][NS] persistent series */
    double *tr = W + (size_t)(P + 1) * NS + tid;  /* [ntr][nth] transient scalars of the current order */
    double *xnew = W + (size_t)(P + 1) * NS + (size_t)a.ntr * nth; /* [dof] */
    double *extra = xnew + a.dof + 8;
#if defined(TB_W2_GENERATED) && TB_W2_LC_SMEM
    /* loop-invariant coupling constants: once per CTA (placement fixed at compile time: shared-memory loads, no generic LD) */
    for (int j = tid; j < a.lc_count; j += nth) extra[j] = a.lc_coef[j];
    const double *lcc = extra;
    extra += a.lc_count;
#elif defined(TB_W2_GENERATED)
    const double *__restrict__ lcc = a.lc_coef;
#else
    const double *lcc = a.lc_coef;
    if (a.lc_smem) { /* loop-invariant coupling constants: once per CTA */
        for (int j = tid; j < a.lc_count; j += nth) extra[j] = a.lc_coef[j];
        lcc = extra;
        extra += a.lc_count;
    }
#endif
    const C2Op *gops = a.gops;
    const double *gnum = a.gnum;
    if (a.g_smem) {
        double *gn = extra;
        for (int j = tid; j < a.ngnum; j += nth) gn[j] = a.gnum[j];
        int *go = (int *)(gn + a.ngnum);
        const int *src = (const int *)a.gops;
        for (int j = tid; j < a.ngops * (int)(sizeof(C2Op) / 4); j += nth) go[j] = src[j];
        gnum = gn;
        gops = (const C2Op *)go;
    }
    if (tid < 64) w2_rinv[tid] = XD(1.0, (double)(tid > 0 ? tid : 1));
    __syncthreads();
#ifdef TB_W2_GENERATED
    double lcreg[TB_W2_LCREG > 0 ? TB_W2_LCREG : 1];
    tb_w2_lcreg_fill(lcreg, lcc, tid);
#endif
    const double inf = __longlong_as_double(0x7FF0000000000000LL), nan = __longlong_as_double(0x7FF8000000000000LL);
    for (;;) {
        if (tid == 0) s_idx = atomicAdd(a.counter, 1ull);
        __syncthreads();
        const long long i = (long long)s_idx;
        __syncthreads();
        if (i >= a.N) break;
        double par[8];
        for (int j = 0; j < a.npar && j < 8; j++) par[j] = a.params_per_traj ? a.params[(long long)j * a.N + i] : a.params[j];
        for (int c = tid; c < a.dof; c += nth) W[a.xsid[c]] = a.q0[(long long)c * a.N + i];
        if (a.tsid >= 0) for (int k = tid; k <= P; k += nth) W[k * NS + a.tsid] = k == 1 ? 1.0 : 0.0;
        double t0 = a.trange ? a.trange[0] : a.t0;
        const double tmax = a.trange ? a.trange[a.nt - 1] : a.tmax;
        const double sign = copysign(1.0, XS(tmax, t0));
        long long nsteps = 1, iter = 2;
        double t1 = a.trange ? a.trange[1] : 0.0;
        int st = 0;
        if (a.trange && a.xv_trange)
            for (int c = tid; c < a.dof; c += nth) {
                a.xv_trange[((long long)c * a.nt) * a.N + i] = a.q0[(long long)c * a.N + i];
                for (long long j = 1; j < a.nt; j++) a.xv_trange[((long long)c * a.nt + j) * a.N + i] = nan;
            }
        __syncthreads();
        while (XM(sign, t0) < XM(sign, tmax)) {
            if (a.tsid >= 0 && tid == 0) W[a.tsid] = t0;
            __syncthreads();
            /* jetcoeffs!: for ord = 0:order-1 (parse_eqs.jl:259-264); every cluster in tape order, then ONE barrier */
            for (int k = 0; k < P; k++) {
#ifdef TB_W2_GENERATED
                tb_w2_groups<FAST>(a, W, tr, par, lcc, gnum, k, tid, lcreg);
#else
                if (a.ngroups > 0) {
                    for (int g = 0; g < a.ngroups; g++)
                        for (int r = tid; r < a.g_count[g]; r += nth)
                            for (int j = 0; j < a.g_len[g]; j++) {
                                C2Op o = gops[a.g_off[g] + j]; /* the same entry for every lane: a broadcast from shared memory */
                                long long no;
                                memcpy(&no, &o.immb, sizeof no);
                                const double lit = ((o.a == -1 && o.para < 0) || (o.b == -1 && o.parb < 0)) ? gnum[no + r] : 0.0;
                                o.imma = lit; o.immb = lit;
                                if (o.flags & 16) o.res += r;
                                if (o.flags & 32) o.a += r;
                                if (o.flags & 64) o.b += r;
                                if (o.flags & 128) o.aux += r;
                                o.lc_off += r;
                                if (o.flags & 256) o.lc_soff += r;
                                exec2<FAST>(a, W, tr, par, lcc, a.lc_smem != 0, o, k);
                            }
                } else
                for (int c = tid; c < a.ncl; c += nth) {
                    const int len = __ldg(a.cl_len + c);
                    for (int j = 0; j < len; j++) exec2<FAST>(a, W, tr, par, lcc, a.lc_smem != 0, a.ops[(size_t)j * a.ncl + c], k);
                }
#endif
                __syncthreads();
            }
            /* stepsize: stepsize.jl:58-99 */
            double h = inf;
            if (a.reltol == 0.0 && a.abstol >= 0.0) {
                /* reltol = 0: every candidate is (abstol / |x_c[k]|)^(1/k), decreasing in |x_c[k]|, so the minimum belongs
                 * to the largest coefficient of each k.  tb_pow errs by < 1 ulp: coefficients more than 2^-40 below the
                 * maximum cannot win after the 1/k-th power; the ones inside that window are evaluated as well.  Same
                 * bits as the loop over all components (the path below), two powers per step instead of 2 dof.
                 * Julia's min propagates NaN (a NaN coefficient, or 0 * Inf from a non-finite |x_c[0]|): tracked apart. */
                double mk0 = 0.0, mk1 = 0.0;
                bool bad = false;
                for (int c = tid; c < a.dof; c += nth) {
                    const double *X = W + a.xsid[c];
                    const double a0 = fabs(X[(P - 1) * NS]), a1 = fabs(X[P * NS]);
                    bad = bad || a0 != a0 || a1 != a1 || ((a0 != 0.0 || a1 != 0.0) && !(fabs(X[0]) < inf));
                    mk0 = fmax(mk0, a0); mk1 = fmax(mk1, a1);
                }
                mk0 = w2_block_reduce(mk0, red, true);
                mk1 = w2_block_reduce(mk1, red, true);
                bad = __syncthreads_or(bad);
                const double th0 = XM(mk0, 0.99999999999909051), th1 = XM(mk1, 0.99999999999909051);
                for (int c = tid; c < a.dof; c += nth) {
                    const double *X = W + a.xsid[c];
                    const double a0 = fabs(X[(P - 1) * NS]), a1 = fabs(X[P * NS]);
                    if (a0 != 0.0 && (a0 == mk0 || a0 > th0)) h = fmin(h, tb_pow(XD(a.abstol, a0), XD(1.0, (double)(P - 1))));
                    if (a1 != 0.0 && (a1 == mk1 || a1 > th1)) h = fmin(h, tb_pow(XD(a.abstol, a1), XD(1.0, (double)P)));
                }
                if (bad) h = nan;
            } else
            for (int c = tid; c < a.dof; c += nth) {
                const double *X = W + a.xsid[c];
                const double x0norm = fabs(X[0]);
                for (int k = P - 1; k <= P; k++) {
                    const double ak = fabs(X[k * NS]);
                    if (ak == 0.0) continue;
                    const double rel = XM(a.reltol, x0norm);
                    const double eps = a.abstol >= rel ? a.abstol : rel;
                    const double hk = tb_pow(XD(eps, ak), XD(1.0, (double)k));
                    h = (h != h || hk != hk) ? nan : fmin(h, hk);
                }
            }
            h = w2_block_reduce(h, red, false);
            if (h == inf) { /* _second_stepsize: stepsize.jl:119-132, max over components */
                double h2 = 0.0;
                for (int c = tid; c < a.dof; c += nth) {
                    const double *X = W + a.xsid[c];
                    bool allzero = true;
                    for (int k = 0; k <= P; k++) if (X[k * NS] != 0.0) { allzero = false; break; }
                    double hc = inf;
                    if (!allzero) {
                        hc = 0.0;
                        for (int k = 1; k <= P - 2; k++) {
                            const double ak = fabs(X[k * NS]);
                            if (ak == 0.0) continue;
                            const double hk = tb_pow(XD(1.0, ak), XD(1.0, (double)k));
                            hc = (hc != hc || hk != hk) ? nan : fmax(hc, hk);
                        }
                    }
                    h2 = (h2 != h2 || hc != hc) ? nan : fmax(h2, hc);
                }
                h = w2_block_reduce(h2, red, true);
            }
            double dt = h < a.minstep ? a.minstep : (h > a.maxstep ? a.maxstep : h);
            if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
            const double rem = XS(tmax, t0);
            const double srem = XM(sign, rem);
            dt = XM(sign, (dt != dt || srem != srem) ? nan : fmin(dt, srem));
            bool fin = true;
            for (int c = tid; c < a.dof; c += nth) {
                const double s = w2_horner(W + a.xsid[c], NS, P, dt);
                xnew[c] = s;
                fin = fin && isfinite(s);
            }
            bool done = false;
            if (a.trange) {
                const double tnext = XA(t0, dt);
                while (XM(sign, t1) < XM(sign, tnext)) {
                    const double d1 = XS(t1, t0);
                    if (a.xv_trange)
                        for (int c = tid; c < a.dof; c += nth)
                            a.xv_trange[((long long)c * a.nt + (iter - 1)) * a.N + i] = w2_horner(W + a.xsid[c], NS, P, d1);
                    iter += 1;
                    t1 = a.trange[iter - 1];
                }
                if (dt == rem) {
                    if (a.xv_trange)
                        for (int c = tid; c < a.dof; c += nth) a.xv_trange[((long long)c * a.nt + (iter - 1)) * a.N + i] = xnew[c];
                    done = true;
                }
                t0 = tnext;
            } else
                t0 = XA(t0, dt);
            __syncthreads();
            for (int c = tid; c < a.dof; c += nth) W[a.xsid[c]] = xnew[c];
            if (done) break;
            nsteps += 1;
            if (nsteps > a.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
            if (__syncthreads_or(!fin)) { st |= TB_ST_NONFINITE; break; }
        }
        __syncthreads();
        if (a.xfinal) for (int c = tid; c < a.dof; c += nth) a.xfinal[(long long)c * a.N + i] = W[a.xsid[c]];
        if (tid == 0) {
            if (a.tfinal) a.tfinal[i] = t0;
            if (a.nsteps) a.nsteps[i] = nsteps - 1;
            if (a.status) a.status[i] = st;
        }
        __syncthreads();
    }
}

#ifndef TB_W2_GENERATED
template <bool FAST> __global__ void __launch_bounds__(256, 1) wide2_kernel(const W2Args a) { wide2_body<FAST>(a); }
#endif

#endif
