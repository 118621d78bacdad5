/*
 * tb_tn.cu -- TaylorN jet transport on sm_100a: taylorinteg with U = TaylorN{Float64}
 * (reference: dispatch of src/integrator/taylorinteg.jl:329-381 / 582-642 on polynomial-valued
 * coefficients; solcoeff! src/TaylorIntegration.jl:54-65; _norminf src/integrator/stepsize.jl:26-42).
 *
 * One CTA integrates one trajectory (persistent CTAs pull trajectory indices from a queue); thread m
 * owns polynomial coefficient m of every series.  All products are GATHERS: the thread sums
 * U[ia]*V[ib] over the precomputed pairs with mono[ia]+mono[ib] = mono[m], in ascending ia -- the
 * order in which TaylorSeries' scatter loops touch slot m (tb_tn_tables.h) -- so no atomics, no
 * write conflicts, and in parity mode the same bits as oracle/tb_tn_oracle.cpp.  Recurrences over
 * homogeneous degree (TaylorN division, power, sincos) advance one degree per barrier.
 * The tape is interpreted (each entry costs thousands of FMAs, dispatch is free).  Jets live in a
 * per-CTA workspace in global memory: (2 dof + ntmp + 1) series x (P+1) x ncoef doubles
 * (pendulum n=2: 60 KB, L1-resident; Kepler n=4 order 8: ~2 MB, L2-resident).
 */
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "tb_internal.h"
#include "tb_math.h"
#include "tb_tn_tables.h"

namespace tb {
void count_launch(long long n);
void set_last_kernel_ms(double ms);
}
struct tb_ode {
    tb::Ode o;
};

namespace {

#define TB_CUDA(call)                                                          \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            tb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
            return TB_ERR_CUDA;                                                \
        }                                                                      \
    } while (0)

struct TnArgs {
    long long N;
    int dof, npar, ntmp, nops, P, nc, d, nser;
    const tb_op_t *ops;
    const unsigned char *deg;     /* [nc] */
    const int *pair_off;          /* [nc+1] */
    const unsigned short *pair_a; /* pairs of slot m, ascending ia */
    const unsigned short *pair_b;
    const unsigned int *pair_ab; /* ia | ib << 16: one load per pair */
    int npairs, pairs_in_smem;
    const double *q0;     /* [dof x nc x N] */
    const double *params; /* [npar x N] or [npar] */
    const double *trange;
    long long nt;
    double t0, tmax, abstol, reltol, minstep, maxstep;
    long long maxsteps;
    int params_per_traj;
    double *work; /* [gridDim.x][nser][P+1][nc] */
    double *xfinal, *tfinal;
    long long *nsteps;
    int *status;
    double *xv_trange; /* [dof x nc x nt x N] */
    unsigned long long *counter;
    const unsigned int *sched; /* execution schedule of the tape: j | 0xffff0000 = entry j alone, j | q << 16 = the pair (j, q) */
    int nsched;
};

template <bool FAST> __device__ __forceinline__ double mad(double a, double b, double c) {
    return FAST ? fma(a, b, c) : __dadd_rn(c, __dmul_rn(a, b));
}

/* per-CTA view of the jets */
struct Jets {
    double *base;
    int P, nc, dof, ntmp;
    __device__ __forceinline__ double *ser(int o) const {
        const int kind = TB_OPERAND_KIND(o), idx = TB_OPERAND_INDEX(o);
        int s = 0;
        if (kind == TB_K_X) s = idx;
        else if (kind == TB_K_DX) s = dof + idx;
        else if (kind == TB_K_TMP) s = 2 * dof + idx;
        else s = 2 * dof + ntmp; /* TB_K_TIME */
        return base + (size_t)s * (P + 1) * nc;
    }
    __device__ __forceinline__ double *co(int o, int k) const { return ser(o) + (size_t)k * nc; }
};

/* acc += sum over the pairs [p0, p1) of slot m of (f*U[ia]) * V[ib]; U and V are polynomials STAGED IN SHARED MEMORY
 * (random 8-byte gathers from the L2-resident workspace cost a 32-byte sector each and made the n = 4 kernel L2-bound) */
template <bool FAST>
__device__ __forceinline__ double gather(double acc, const TnArgs &a, int p0, int p1, const double *__restrict__ U,
                                         const double *__restrict__ V, bool hasf, double f) {
    const unsigned int *__restrict__ pab = a.pair_ab;
    if (hasf) {
        for (int p = p0; p < p1; p++) {
            const unsigned int ab = pab[p];
            acc = mad<FAST>(__dmul_rn(f, U[ab & 0xffffu]), V[ab >> 16], acc);
        }
    } else {
#pragma unroll 4
        for (int p = p0; p < p1; p++) {
            const unsigned int ab = pab[p];
            acc = mad<FAST>(U[ab & 0xffffu], V[ab >> 16], acc);
        }
    }
    return acc;
}

/* Time-index sum of polynomial products, sum_{i = ilo..ihi} (f_i U_i) * V_i, operands staged through shared memory:
 * while the CTA gathers from one pair of buffers every thread copies ITS coefficient of the next operands into the
 * other pair (coalesced reads of the workspace), one barrier per i.  term(i, U, V, hasf, f) names the operands.    */
template <bool FAST, class F>
__device__ __forceinline__ double staged_sum(double acc, const TnArgs &a, double *__restrict__ stg, int m, bool act, int p0, int p1,
                                             int ilo, int ihi, F term) {
    const int nc = a.nc;
    const double *U, *V;
    bool hasf; double f;
    int buf = 0;
    if (ilo <= ihi) {
        term(ilo, U, V, hasf, f);
        if (act) { stg[m] = U[m]; stg[nc + m] = V[m]; }
    }
    __syncthreads();
    for (int i = ilo; i <= ihi; i++) {
        term(i, U, V, hasf, f);
        const bool hf = hasf; const double ff = f;
        if (i < ihi) {
            term(i + 1, U, V, hasf, f);
            if (act) { stg[(2 * (buf ^ 1)) * nc + m] = U[m]; stg[(2 * (buf ^ 1) + 1) * nc + m] = V[m]; }
        }
        if (act) acc = gather<FAST>(acc, a, p0, p1, stg + (2 * buf) * nc, stg + (2 * buf + 1) * nc, hf, ff);
        __syncthreads();
        buf ^= 1;
    }
    return acc;
}

/* sincos! at order k >= 1: ts = sum_i (i A_i) C_{k-i} and tc = sum_i (i A_i) S_{k-i} share the operand (i A_i): ONE
 * staging per time index (three polynomials per stage, double buffered: the six buffers), one barrier instead of
 * two, one gather of the shared operand, and the two dependent add chains interleaved.  Each sum keeps its order. */
template <bool FAST>
__device__ __forceinline__ void staged_sincos(double &ts, double &tc, const TnArgs &a, const Jets &J, double *__restrict__ stg,
                                              int m, bool act, int p0, int p1, int oa, int ocos, int osin, int k) {
    const int nc = a.nc;
    const unsigned int *__restrict__ pab = a.pair_ab;
    int buf = 0;
    if (act) { stg[m] = J.co(oa, 1)[m]; stg[nc + m] = J.co(ocos, k - 1)[m]; stg[2 * nc + m] = J.co(osin, k - 1)[m]; }
    __syncthreads();
    for (int i = 1; i <= k; i++) {
        if (i < k && act) {
            double *d = stg + (3 * (buf ^ 1)) * nc;
            d[m] = J.co(oa, i + 1)[m]; d[nc + m] = J.co(ocos, k - i - 1)[m]; d[2 * nc + m] = J.co(osin, k - i - 1)[m];
        }
        if (act) {
            const double *sU = stg + (3 * buf) * nc, *sVc = sU + nc, *sVs = sU + 2 * nc;
            const bool hf = i >= 2;
            const double f = (double)i;
#pragma unroll 2
            for (int p = p0; p < p1; p++) {
                const unsigned int ab = pab[p];
                const double u0 = sU[ab & 0xffffu];
                const double u = hf ? __dmul_rn(f, u0) : u0;
                ts = mad<FAST>(u, sVc[ab >> 16], ts);
                tc = mad<FAST>(u, sVs[ab >> 16], tc);
            }
        }
        __syncthreads();
        buf ^= 1;
    }
}

/* Q = Nn / B over homogeneous degree; nn = Nn[m] in a register.  Q_g = (Nn_g - sum_{j<g} Q_j B_{g-j}) / B_0 */
template <bool FAST>
__device__ void tn_div(const TnArgs &a, int m, int degm, double *Q, double nn, const double *B, double *__restrict__ stg) {
    const bool act = m < a.nc;
    double *sQ = stg, *sB = stg + a.nc;
    if (act) sB[m] = B[m];
    __syncthreads();
    const int p0 = act ? a.pair_off[m] : 0, p1 = act ? a.pair_off[m + 1] - 1 : 0; /* last pair is (m, 0) */
    for (int g = 0; g <= a.d; g++) {
        if (act && degm == g) {
            const double t = gather<FAST>(0.0, a, p0, p1, sQ, sB, false, 0.0);
            const double q = __ddiv_rn(__dsub_rn(nn, t), sB[0]);
            sQ[m] = q;
            Q[m] = q;
        }
        __syncthreads();
    }
}

/* C = A^r: C_0 = A_0^r; C_g = sum_{j<g} ((r(g-j) - j) C_j) A_{g-j} / (g A_0) */
template <bool FAST>
__device__ void tn_pow(const TnArgs &a, int m, int degm, double *C, const double *A, double r, double *__restrict__ stg) {
    const bool act = m < a.nc;
    double *sC = stg, *sA = stg + a.nc;
    if (act) sA[m] = A[m];
    __syncthreads();
    if (m == 0) { const double c0 = tb_pow(sA[0], r); sC[0] = c0; C[0] = c0; }
    __syncthreads();
    const int p0 = act ? a.pair_off[m] : 0, p1 = act ? a.pair_off[m + 1] - 1 : 0;
    for (int g = 1; g <= a.d; g++) {
        if (act && degm == g) {
            double t = 0.0;
            for (int p = p0; p < p1; p++) {
                const unsigned int ab = a.pair_ab[p];
                const int ia = ab & 0xffffu, ib = ab >> 16;
                const int j = a.deg[ia];
                const double f = __dsub_rn(__dmul_rn(r, (double)(g - j)), (double)j);
                t = mad<FAST>(__dmul_rn(f, sC[ia]), sA[ib], t);
            }
            const double c = __ddiv_rn(t, __dmul_rn((double)g, sA[0]));
            sC[m] = c;
            C[m] = c;
        }
        __syncthreads();
    }
}

/* S = sin A, Cc = cos A: S_g = sum_{j>=1} (j A_j) Cc_{g-j} / g, Cc_g = -sum_{j>=1} (j A_j) S_{g-j} / g */
template <bool FAST>
__device__ void tn_sincos(const TnArgs &a, int m, int degm, double *S, double *Cc, const double *A, double *__restrict__ stg) {
    const bool act = m < a.nc;
    double *sS = stg, *sC = stg + a.nc, *sA = stg + 2 * a.nc;
    if (act) sA[m] = A[m];
    __syncthreads();
    if (m == 0) { double s, c; tb_sincos(sA[0], &s, &c); sS[0] = s; sC[0] = c; S[0] = s; Cc[0] = c; }
    __syncthreads();
    const int p0 = act ? a.pair_off[m] + 1 : 0, p1 = act ? a.pair_off[m + 1] : 0; /* first pair is (0, m) */
    for (int g = 1; g <= a.d; g++) {
        if (act && degm == g) {
            double ts = 0.0, tc = 0.0;
            for (int p = p0; p < p1; p++) {
                const unsigned int ab = a.pair_ab[p];
                const int ia = ab & 0xffffu, ib = ab >> 16;
                const int j = a.deg[ia];
                double u = sA[ia];
                if (j >= 2) u = __dmul_rn((double)j, u);
                ts = mad<FAST>(u, sC[ib], ts);
                tc = mad<FAST>(u, sS[ib], tc);
            }
            const double sv = __ddiv_rn(ts, (double)g), cv = __ddiv_rn(-tc, (double)g);
            sS[m] = sv; sC[m] = cv;
            S[m] = sv; Cc[m] = cv;
        }
        __syncthreads();
    }
}

/* one tape entry at time order k; every thread of the CTA calls it; ends with a barrier */
template <bool FAST>
__device__ void run_op(const TnArgs &a, const Jets &J, const double *par, double *scratch, double *stg, const tb_op_t &op, int k,
                       int m, int degm) {
    const int ka = TB_OPERAND_KIND(op.a), kb = TB_OPERAND_KIND(op.b);
    const bool sa = ka == TB_K_PAR || ka == TB_K_IMM, sb = kb == TB_K_PAR || kb == TB_K_IMM;
    const double na = ka == TB_K_PAR ? par[TB_OPERAND_INDEX(op.a)] : op.imm;
    const double nb = kb == TB_K_PAR ? par[TB_OPERAND_INDEX(op.b)] : op.imm;
    const bool act = m < a.nc;
    double *R = J.co(op.res, k);
    const int p0 = act ? a.pair_off[m] : 0, p1 = act ? a.pair_off[m + 1] : 0;
    int opc = op.op;
    if (opc == TB_OP_POW) {
        if (op.imm == 1.0) opc = TB_OP_IDENTITY;
        else if (op.imm == 2.0) opc = TB_OP_SQR;
    }
    switch (opc) {
        case TB_OP_IDENTITY:
            if (act) R[m] = sa ? ((k == 0 && m == 0) ? na : 0.0) : J.co(op.a, k)[m];
            break;
        case TB_OP_NEG:
            if (act) R[m] = -J.co(op.a, k)[m];
            break;
        case TB_OP_ADD:
        case TB_OP_SUB:
            if (act) {
                const bool add = opc == TB_OP_ADD, c0 = k == 0 && m == 0;
                double v;
                if (sa && sb) v = c0 ? (add ? __dadd_rn(na, nb) : __dsub_rn(na, nb)) : 0.0;
                else if (sa) { const double b = J.co(op.b, k)[m]; v = c0 ? (add ? __dadd_rn(na, b) : __dsub_rn(na, b)) : (add ? b : -b); }
                else if (sb) { const double x = J.co(op.a, k)[m]; v = c0 ? (add ? __dadd_rn(x, nb) : __dsub_rn(x, nb)) : x; }
                else { const double x = J.co(op.a, k)[m], b = J.co(op.b, k)[m]; v = add ? __dadd_rn(x, b) : __dsub_rn(x, b); }
                R[m] = v;
            }
            break;
        case TB_OP_MUL:
            if (sa && sb) { if (act) R[m] = (k == 0 && m == 0) ? __dmul_rn(na, nb) : 0.0; }
            else if (sa || sb) { if (act) R[m] = __dmul_rn(sa ? na : nb, J.co(sa ? op.b : op.a, k)[m]); }
            else {
                const double acc = staged_sum<FAST>(0.0, a, stg, m, act, p0, p1, 0, k,
                    [&](int i, const double *&U, const double *&V, bool &hf, double &f) { U = J.co(op.a, i); V = J.co(op.b, k - i); hf = false; f = 0.0; });
                if (act) R[m] = acc;
            }
            break;
        case TB_OP_SQR: {
            const int kodd = k & 1, kend = k == 0 ? 0 : (k - 2 + kodd) / 2;
            double acc = staged_sum<FAST>(0.0, a, stg, m, act, p0, p1, 0, kend,
                [&](int i, const double *&U, const double *&V, bool &hf, double &f) { U = J.co(op.a, i); V = J.co(op.a, k - i); hf = false; f = 0.0; });
            if (k > 0) {
                acc = __dmul_rn(2.0, acc);
                if (!kodd) acc = staged_sum<FAST>(acc, a, stg, m, act, p0, p1, k / 2, k / 2,
                    [&](int i, const double *&U, const double *&V, bool &hf, double &f) { U = J.co(op.a, i); V = J.co(op.a, i); hf = false; f = 0.0; });
            }
            if (act) R[m] = acc;
            break;
        }
        case TB_OP_DIV:
            if (sb) {
                if (act) R[m] = sa ? ((k == 0 && m == 0) ? __ddiv_rn(na, nb) : 0.0) : __ddiv_rn(J.co(op.a, k)[m], nb);
            } else {
                const double t = staged_sum<FAST>(0.0, a, stg, m, act, p0, p1, 0, k - 1,
                    [&](int i, const double *&U, const double *&V, bool &hf, double &f) { U = J.co(op.res, i); V = J.co(op.b, k - i); hf = false; f = 0.0; });
                double nn = 0.0;
                if (act) {
                    const double num = sa ? ((k == 0 && m == 0) ? na : 0.0) : J.co(op.a, k)[m];
                    nn = __dsub_rn(num, t);
                }
                tn_div<FAST>(a, m, degm, R, nn, J.co(op.b, 0), stg);
            }
            break;
        case TB_OP_POW:
            if (k == 0) tn_pow<FAST>(a, m, degm, R, J.co(op.a, 0), op.imm, stg);
            else {
                const double t = staged_sum<FAST>(0.0, a, stg, m, act, p0, p1, 0, k - 1,
                    [&](int i, const double *&U, const double *&V, bool &hf, double &f) {
                        U = J.co(op.res, i); V = J.co(op.a, k - i); hf = true;
                        f = __dsub_rn(__dmul_rn(op.imm, (double)(k - i)), (double)i);
                    });
                if (act) scratch[m] = __dmul_rn((double)k, J.co(op.a, 0)[m]);
                __syncthreads();
                tn_div<FAST>(a, m, degm, R, t, scratch, stg);
            }
            break;
        case TB_OP_SINCOS: {
            double *Cc = J.co(op.aux, k);
            if (k == 0) tn_sincos<FAST>(a, m, degm, R, Cc, J.co(op.a, 0), stg);
            else {
                double ts = 0.0, tc = 0.0;
                staged_sincos<FAST>(ts, tc, a, J, stg, m, act, p0, p1, op.a, op.aux, op.res, k);
                if (act) { R[m] = __ddiv_rn(ts, (double)k); Cc[m] = __ddiv_rn(-tc, (double)k); }
            }
            break;
        }
        default: break; /* rejected on the host */
    }
    __syncthreads();
}

/* ---- two independent tape entries at once ------------------------------------------------------------------
 * Entries whose time-index sums do not depend on each other (x^2 and y^2, x / r^3 and y / r^3 of kepler1!) run as a
 * PAIR (host schedule, tn_schedule): one staging of the four operand polynomials per time index, one barrier instead
 * of two, and the two dependent add chains of a thread interleaved (the same index pairs serve both).  Each sum keeps
 * its own order and its own finalisation (doubling / division over the homogeneous degrees): same bits.            */
struct SumDesc { bool has; int oU, oV, ilo, ihi, fmode; double r; };

__device__ __forceinline__ SumDesc sum_desc(const tb_op_t &op, int k) {
    SumDesc d{false, 0, 0, 0, -1, 0, 0.0};
    const int ka = TB_OPERAND_KIND(op.a), kb = TB_OPERAND_KIND(op.b);
    const bool sa = ka == TB_K_PAR || ka == TB_K_IMM, sb = kb == TB_K_PAR || kb == TB_K_IMM;
    int opc = op.op;
    if (opc == TB_OP_POW) {
        if (op.imm == 1.0) opc = TB_OP_IDENTITY;
        else if (op.imm == 2.0) opc = TB_OP_SQR;
    }
    if (opc == TB_OP_MUL && !sa && !sb) { d.has = true; d.oU = op.a; d.oV = op.b; d.ilo = 0; d.ihi = k; }
    else if (opc == TB_OP_SQR) { const int kodd = k & 1; d.has = true; d.oU = op.a; d.oV = op.a; d.ilo = 0; d.ihi = k == 0 ? 0 : (k - 2 + kodd) / 2; }
    else if (opc == TB_OP_DIV && !sb) { d.has = true; d.oU = op.res; d.oV = op.b; d.ilo = 0; d.ihi = k - 1; }
    else if (opc == TB_OP_POW && k > 0) { d.has = true; d.oU = op.res; d.oV = op.a; d.ilo = 0; d.ihi = k - 1; d.fmode = 1; d.r = op.imm; }
    return d;
}

/* what run_op does after its time-index sum `acc`; ends with a barrier */
template <bool FAST>
__device__ void finish_op(const TnArgs &a, const Jets &J, const double *par, double *scratch, double *stg, const tb_op_t &op, int k,
                          int m, int degm, double acc) {
    const int ka = TB_OPERAND_KIND(op.a);
    const bool sa = ka == TB_K_PAR || ka == TB_K_IMM;
    const double na = ka == TB_K_PAR ? par[TB_OPERAND_INDEX(op.a)] : op.imm;
    const bool act = m < a.nc;
    double *R = J.co(op.res, k);
    const int p0 = act ? a.pair_off[m] : 0, p1 = act ? a.pair_off[m + 1] : 0;
    int opc = op.op;
    if (opc == TB_OP_POW && op.imm == 2.0) opc = TB_OP_SQR;
    if (opc == TB_OP_MUL) { if (act) R[m] = acc; }
    else if (opc == TB_OP_SQR) {
        if (k > 0) {
            acc = __dmul_rn(2.0, acc);
            if (!(k & 1)) acc = staged_sum<FAST>(acc, a, stg, m, act, p0, p1, k / 2, k / 2,
                [&](int i, const double *&U, const double *&V, bool &hf, double &f) { U = J.co(op.a, i); V = J.co(op.a, i); hf = false; f = 0.0; });
        }
        if (act) R[m] = acc;
    } else if (opc == TB_OP_DIV) {
        double nn = 0.0;
        if (act) {
            const double num = sa ? ((k == 0 && m == 0) ? na : 0.0) : J.co(op.a, k)[m];
            nn = __dsub_rn(num, acc);
        }
        tn_div<FAST>(a, m, degm, R, nn, J.co(op.b, 0), stg);
    } else { /* TB_OP_POW, k > 0 */
        if (act) scratch[m] = __dmul_rn((double)k, J.co(op.a, 0)[m]);
        __syncthreads();
        tn_div<FAST>(a, m, degm, R, acc, scratch, stg);
    }
    __syncthreads();
}

template <bool FAST>
__device__ void run_pair(const TnArgs &a, const Jets &J, const double *par, double *scratch, double *stg, const tb_op_t &o1,
                         const tb_op_t &o2, int k, int m, int degm) {
    const SumDesc d1 = sum_desc(o1, k), d2 = sum_desc(o2, k);
    if (!d1.has || !d2.has) { /* order 0 of a general power: no sum to share */
        run_op<FAST>(a, J, par, scratch, stg, o1, k, m, degm);
        run_op<FAST>(a, J, par, scratch, stg, o2, k, m, degm);
        return;
    }
    const int nc = a.nc;
    const bool act = m < nc;
    const int p0 = act ? a.pair_off[m] : 0, p1 = act ? a.pair_off[m + 1] : 0;
    const unsigned int *__restrict__ pab = a.pair_ab;
    const int lo = min(d1.ilo, d2.ilo), hi = max(d1.ihi, d2.ihi);
    double acc1 = 0.0, acc2 = 0.0;
    auto stage = [&](int i, int buf) {
        if (!act) return;
        double *d = stg + (4 * buf) * nc;
        if (i >= d1.ilo && i <= d1.ihi) { d[m] = J.co(d1.oU, i)[m]; d[nc + m] = J.co(d1.oV, k - i)[m]; }
        if (i >= d2.ilo && i <= d2.ihi) { d[2 * nc + m] = J.co(d2.oU, i)[m]; d[3 * nc + m] = J.co(d2.oV, k - i)[m]; }
    };
    int buf = 0;
    if (lo <= hi) stage(lo, 0);
    __syncthreads();
    for (int i = lo; i <= hi; i++) {
        if (i < hi) stage(i + 1, buf ^ 1);
        if (act) {
            const double *U1 = stg + (4 * buf) * nc, *V1 = U1 + nc, *U2 = U1 + 2 * nc, *V2 = U1 + 3 * nc;
            const bool in1 = i >= d1.ilo && i <= d1.ihi, in2 = i >= d2.ilo && i <= d2.ihi;
            const double f1 = d1.fmode ? __dsub_rn(__dmul_rn(d1.r, (double)(k - i)), (double)i) : 0.0;
            const double f2 = d2.fmode ? __dsub_rn(__dmul_rn(d2.r, (double)(k - i)), (double)i) : 0.0;
            if (in1 && in2) {
#pragma unroll 2
                for (int p = p0; p < p1; p++) {
                    const unsigned int ab = pab[p];
                    const int ia = ab & 0xffffu, ib = ab >> 16;
                    const double u1 = U1[ia], u2 = U2[ia];
                    acc1 = mad<FAST>(d1.fmode ? __dmul_rn(f1, u1) : u1, V1[ib], acc1);
                    acc2 = mad<FAST>(d2.fmode ? __dmul_rn(f2, u2) : u2, V2[ib], acc2);
                }
            } else if (in1) {
                for (int p = p0; p < p1; p++) {
                    const unsigned int ab = pab[p];
                    const double u1 = U1[ab & 0xffffu];
                    acc1 = mad<FAST>(d1.fmode ? __dmul_rn(f1, u1) : u1, V1[ab >> 16], acc1);
                }
            } else if (in2) {
                for (int p = p0; p < p1; p++) {
                    const unsigned int ab = pab[p];
                    const double u2 = U2[ab & 0xffffu];
                    acc2 = mad<FAST>(d2.fmode ? __dmul_rn(f2, u2) : u2, V2[ab >> 16], acc2);
                }
            }
        }
        __syncthreads();
        buf ^= 1;
    }
    finish_op<FAST>(a, J, par, scratch, stg, o1, k, m, degm, acc1);
    finish_op<FAST>(a, J, par, scratch, stg, o2, k, m, degm, acc2);
}

/* block-wide max of |v| (order-independent, so bit-identical to the oracle's sequential _norminf) */
__device__ double block_max(double v, double *red) {
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < nw; i++) r = fmax(r, red[i]);
    return r;
}

template <bool FAST> __global__ void __launch_bounds__(1024) tn_kernel(const TnArgs a_in) {
    __shared__ double red[32];
    __shared__ unsigned long long s_idx;
    extern __shared__ double scratch[]; /* [nc] + staging buffers [8][nc] (+ the pair table) */
    TnArgs a = a_in;
    double *stg = scratch + a.nc;
    if (a.pairs_in_smem) { /* the index pairs are read once per product term: keep them next to the operands */
        unsigned int *sp = (unsigned int *)(scratch + 9 * a.nc);
        for (int q = threadIdx.x; q < a.npairs; q += blockDim.x) sp[q] = a_in.pair_ab[q];
        a.pair_ab = sp;
        __syncthreads();
    }
    const int m = threadIdx.x, nc = a.nc, P = a.P;
    const bool act = m < nc;
    const int degm = act ? a.deg[m] : -1;
    Jets J{a.work + (size_t)blockIdx.x * a.nser * (P + 1) * nc, P, nc, a.dof, a.ntmp};
    const int TIME = TB_OPERAND(TB_K_TIME, 0);
    const double inf = __longlong_as_double(0x7FF0000000000000LL), nan = __longlong_as_double(0x7FF8000000000000LL);
    for (;;) {
        if (m == 0) s_idx = atomicAdd(a.counter, 1ull);
        __syncthreads();
        const long long i = (long long)s_idx;
        __syncthreads();
        if (i >= a.N) break;
        double par[8];
        for (int j = 0; j < a.npar && j < 8; j++) par[j] = a.params_per_traj ? a.params[(long long)j * a.N + i] : a.params[j];
        /* init_expansions / update_cache!: cache.jl:176-192, 435-442 */
        if (act) {
            for (int c = 0; c < a.dof; c++) J.co(TB_OPERAND(TB_K_X, c), 0)[m] = a.q0[((long long)c * nc + m) * a.N + i];
            for (int k = 0; k <= P; k++) J.co(TIME, k)[m] = (k == 1 && m == 0) ? 1.0 : 0.0;
        }
        double t0 = a.trange ? a.trange[0] : a.t0;
        const double tmax = a.trange ? a.trange[a.nt - 1] : a.tmax;
        const double sign = copysign(1.0, __dsub_rn(tmax, t0));
        long long nsteps = 1, iter = 2;
        double t1 = a.trange ? a.trange[1] : 0.0;
        int st = 0;
        if (a.trange && a.xv_trange && act)
            for (int c = 0; c < a.dof; c++) {
                a.xv_trange[(((long long)c * nc + m) * a.nt) * a.N + i] = J.co(TB_OPERAND(TB_K_X, c), 0)[m];
                for (long long j = 1; j < a.nt; j++) a.xv_trange[(((long long)c * nc + m) * a.nt + j) * a.N + i] = nan;
            }
        __syncthreads();
        while (__dmul_rn(sign, t0) < __dmul_rn(sign, tmax)) {
            /* jetcoeffs!: parse_eqs.jl:259-264, solcoeff! TaylorIntegration.jl:54-65 */
            if (m == 0) J.co(TIME, 0)[0] = t0;
            __syncthreads();
            for (int k = 0; k < P; k++) {
                for (int s = 0; s < a.nsched; s++) {
                    const unsigned int e = a.sched[s];
                    const int j = e & 0xffffu, q = e >> 16;
                    if (q == 0xffff) run_op<FAST>(a, J, par, scratch, stg, a.ops[j], k, m, degm);
                    else run_pair<FAST>(a, J, par, scratch, stg, a.ops[j], a.ops[q], k, m, degm);
                }
                if (act)
                    for (int c = 0; c < a.dof; c++)
                        J.co(TB_OPERAND(TB_K_X, c), k + 1)[m] = __ddiv_rn(J.co(TB_OPERAND(TB_K_DX, c), k)[m], (double)(k + 1));
                __syncthreads();
            }
            /* stepsize: stepsize.jl:58-99 with _norminf (26-42) = max |coefficient| */
            double h = inf;
            for (int c = 0; c < a.dof; c++) {
                const int X = TB_OPERAND(TB_K_X, c);
                const double x0norm = block_max(act ? fabs(J.co(X, 0)[m]) : 0.0, red);
                for (int k = P - 1; k <= P; k++) {
                    const double nk = block_max(act ? fabs(J.co(X, k)[m]) : 0.0, red);
                    if (nk == 0.0) continue;
                    const double rel = __dmul_rn(a.reltol, x0norm);
                    const double eps = a.abstol >= rel ? a.abstol : rel;
                    h = fmin(h, tb_pow(__ddiv_rn(eps, nk), __ddiv_rn(1.0, (double)k)));
                }
            }
            if (h == inf) { /* _second_stepsize: stepsize.jl:119-132 */
                h = 0.0;
                for (int c = 0; c < a.dof; c++) {
                    const int X = TB_OPERAND(TB_K_X, c);
                    double hc = 0.0;
                    bool allzero = block_max(act ? fabs(J.co(X, 0)[m]) : 0.0, red) == 0.0 &&
                                   block_max(act ? fabs(J.co(X, P - 1)[m]) : 0.0, red) == 0.0 &&
                                   block_max(act ? fabs(J.co(X, P)[m]) : 0.0, red) == 0.0;
                    for (int k = 1; k <= P - 2; k++) {
                        const double nk = block_max(act ? fabs(J.co(X, k)[m]) : 0.0, red);
                        if (nk == 0.0) continue;
                        allzero = false;
                        hc = fmax(hc, tb_pow(__ddiv_rn(1.0, nk), __ddiv_rn(1.0, (double)k)));
                    }
                    h = fmax(h, allzero ? inf : hc);
                }
            }
            double dt = h < a.minstep ? a.minstep : (h > a.maxstep ? a.maxstep : h);
            if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
            const double rem = __dsub_rn(tmax, t0);
            dt = __dmul_rn(sign, fmin(dt, __dmul_rn(sign, rem)));
            double xn[16];
            bool fin = true;
            if (act)
                for (int c = 0; c < a.dof; c++) {
                    const double *X = J.ser(TB_OPERAND(TB_K_X, c));
                    double s = X[(size_t)P * nc + m];
                    for (int k = P - 1; k >= 0; k--) s = __dadd_rn(__dmul_rn(s, dt), X[(size_t)k * nc + m]);
                    xn[c] = s;
                    fin = fin && isfinite(s);
                }
            bool done = false;
            if (a.trange) {
                const double tnext = __dadd_rn(t0, dt);
                while (__dmul_rn(sign, t1) < __dmul_rn(sign, tnext)) {
                    const double d1 = __dsub_rn(t1, t0);
                    if (act && a.xv_trange)
                        for (int c = 0; c < a.dof; c++) {
                            const double *X = J.ser(TB_OPERAND(TB_K_X, c));
                            double s = X[(size_t)P * nc + m];
                            for (int k = P - 1; k >= 0; k--) s = __dadd_rn(__dmul_rn(s, d1), X[(size_t)k * nc + m]);
                            a.xv_trange[(((long long)c * nc + m) * a.nt + (iter - 1)) * a.N + i] = s;
                        }
                    iter += 1;
                    t1 = a.trange[iter - 1];
                }
                if (dt == rem) {
                    if (act && a.xv_trange)
                        for (int c = 0; c < a.dof; c++) a.xv_trange[(((long long)c * nc + m) * a.nt + (iter - 1)) * a.N + i] = xn[c];
                    done = true;
                }
                t0 = tnext;
            } else
                t0 = __dadd_rn(t0, dt);
            __syncthreads(); /* every thread has read the old jets */
            if (act)
                for (int c = 0; c < a.dof; c++) J.co(TB_OPERAND(TB_K_X, c), 0)[m] = xn[c];
            if (done) break;
            nsteps += 1;
            if (nsteps > a.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
            if (__syncthreads_or(!fin)) { st |= TB_ST_NONFINITE; break; }
        }
        __syncthreads();
        if (act && a.xfinal)
            for (int c = 0; c < a.dof; c++) a.xfinal[((long long)c * nc + m) * a.N + i] = J.co(TB_OPERAND(TB_K_X, c), 0)[m];
        if (m == 0) {
            if (a.tfinal) a.tfinal[i] = t0;
            if (a.nsteps) a.nsteps[i] = nsteps - 1;
            if (a.status) a.status[i] = st;
        }
        __syncthreads();
    }
}

} // namespace

/* ---- host side ------------------------------------------------------------------------------------ */
struct tb_tn {
    tb::Ode ode;
    int order = 0, nvars = 0, varorder = 0, mode = 0;
    tbtn::Tables T;
    /* device copies (current device of the first integrate call) */
    int device = -1;
    tb_op_t *d_ops = nullptr;
    unsigned char *d_deg = nullptr;
    int *d_off = nullptr;
    unsigned short *d_pa = nullptr, *d_pb = nullptr;
    unsigned int *d_pab = nullptr;
    unsigned int *d_sched = nullptr;
    std::vector<unsigned int> sched; /* execution schedule (tn_schedule) */
    unsigned long long serial = 0; /* identity of this handle (tb_tn_integrate_static: whose tables sit in the static array) */
};

/* Execution schedule of the tape: entries in tape order, except that an entry with a time-index sum (series x series
 * product, square, power, series / series) is PAIRED with the next such entry that does not read its result; the
 * entries between the two move in front of the pair (they cannot depend on the first either, or the pair is not
 * formed).  Temporaries and dx are written once (the tracer is SSA), so moving independent entries is safe.
 * TB_TN_NO_PAIRS=1 keeps the plain tape order.                                                                    */
static void tn_schedule(tb_tn *t) {
    const tb_ode_desc_t &d = t->ode.desc;
    const std::vector<tb_op_t> &ops = t->ode.ops;
    auto numeric = [](int32_t o) { const int k = TB_OPERAND_KIND(o); return k == TB_K_PAR || k == TB_K_IMM; };
    auto summing = [&](const tb_op_t &o) {
        if (o.op == TB_OP_MUL) return !numeric(o.a) && !numeric(o.b);
        if (o.op == TB_OP_SQR) return true;
        if (o.op == TB_OP_POW) return o.imm != 1.0;
        if (o.op == TB_OP_DIV) return !numeric(o.b);
        return false;
    };
    auto reads = [&](const tb_op_t &o, int32_t loc) { return o.a == loc || o.b == loc; };
    std::vector<char> done(d.nops, 0);
    t->sched.clear();
    const bool pairs = !getenv("TB_TN_NO_PAIRS");
    for (int j = 0; j < d.nops; j++) {
        if (done[j]) continue;
        int q = -1;
        if (pairs && summing(ops[j])) {
            bool blocked = false;
            for (int c = j + 1; c < d.nops && !blocked; c++) {
                if (done[c]) continue;
                const bool dep = reads(ops[c], ops[j].res) || (ops[j].op == TB_OP_SINCOS && reads(ops[c], ops[j].aux));
                if (dep) { blocked = true; break; } /* everything after a dependent entry stays behind the pair */
                if (summing(ops[c])) { q = c; break; }
                if (ops[c].op == TB_OP_SINCOS) { blocked = true; break; } /* keep the (internally fused) sincos! in place */
            }
        }
        if (q >= 0) {
            for (int c = j + 1; c < q; c++)
                if (!done[c]) { t->sched.push_back((unsigned int)c | 0xffff0000u); done[c] = 1; }
            t->sched.push_back((unsigned int)j | ((unsigned int)q << 16));
            done[j] = done[q] = 1;
        } else {
            t->sched.push_back((unsigned int)j | 0xffff0000u);
            done[j] = 1;
        }
    }
}

extern "C" {

tb_tn_t *tb_tn_create(const tb_ode_t *ode, int order, int nvars, int varorder, int mode) {
    if (!ode || order < 1 || order > 60 || nvars < 1 || nvars > 8 || varorder < 0 || varorder > 20) {
        tb::set_error("tb_tn_create: bad argument");
        return nullptr;
    }
    const tb_ode_desc_t &d = ode->o.desc;
    if (d.dof > 16 || d.npar > 8) { tb::set_error("tb_tn_create: dof <= 16 and npar <= 8 supported"); return nullptr; }
    for (int j = 0; j < d.nops; j++) {
        const tb_op_t &op = ode->o.ops[j];
        const int ka = TB_OPERAND_KIND(op.a);
        const bool sa = ka == TB_K_PAR || ka == TB_K_IMM;
        bool ok = false;
        switch (op.op) {
            case TB_OP_IDENTITY: case TB_OP_ADD: case TB_OP_SUB: case TB_OP_MUL: case TB_OP_DIV: ok = true; break;
            case TB_OP_NEG: case TB_OP_SQR: case TB_OP_POW: case TB_OP_SINCOS: ok = !sa; break;
            default: break;
        }
        if (!ok) { tb::set_error("tb_tn_create: tape entry " + std::to_string(j) + " has no TaylorN kernel yet (supported: identity, neg, add, sub, mul, div, sqr, pow, sincos)"); return nullptr; }
    }
    tb_tn *t = new tb_tn;
    t->ode = ode->o;
    t->ode.desc.ops = t->ode.ops.data();
    t->order = order; t->nvars = nvars; t->varorder = varorder; t->mode = mode;
    t->T = tbtn::make_tables(nvars, varorder);
    if (t->T.ncoef > 1024) { tb::set_error("tb_tn_create: more than 1024 polynomial coefficients"); delete t; return nullptr; }
    tn_schedule(t);
    static std::atomic<unsigned long long> next_serial{1};
    t->serial = next_serial++;
    return t;
}

void tb_tn_free(tb_tn_t *t) {
    if (!t) return;
    if (t->d_ops) { cudaFree(t->d_ops); cudaFree(t->d_deg); cudaFree(t->d_off); cudaFree(t->d_pa); cudaFree(t->d_pb); cudaFree(t->d_pab); cudaFree(t->d_sched); }
    delete t;
}

int tb_tn_ncoef(const tb_tn_t *t) { return t ? t->T.ncoef : TB_ERR_ARG; }

/* execution schedule of the tape (tn_schedule): entry = j | 0xffff0000 (tape entry j alone) or j | q << 16 (the
 * entries j and q as a pair); returns the number of entries, writes at most `cap` of them */
int tb_tn_schedule(const tb_tn_t *t, uint32_t *sched, int cap) {
    if (!t) return TB_ERR_ARG;
    for (int i = 0; i < cap && i < (int)t->sched.size(); i++) sched[i] = t->sched[i];
    return (int)t->sched.size();
}

int tb_tn_monomials(const tb_tn_t *t, int32_t *expo) {
    if (!t || !expo) return TB_ERR_ARG;
    for (size_t i = 0; i < t->T.expo.size(); i++) expo[i] = t->T.expo[i];
    return t->T.ncoef;
}

int tb_tn_integrate(tb_tn_t *t, int64_t N, const double *q0, const double *params, double t0, double tmax,
                    const double *trange, int64_t nt, const tb_opts_t *o, const tb_tn_out_t *out) {
    if (!t || !q0 || !o || !out || N < 0 || (trange && nt < 2)) { tb::set_error("tb_tn_integrate: bad argument"); return TB_ERR_ARG; }
    if (N == 0) return TB_OK;
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    const tb_ode_desc_t &d = t->ode.desc;
    const int nc = t->T.ncoef, P = t->order;
    if (!t->d_ops || t->device != dev) {
        t->device = dev;
        TB_CUDA(cudaMalloc(&t->d_ops, sizeof(tb_op_t) * (d.nops > 0 ? d.nops : 1)));
        TB_CUDA(cudaMemcpy(t->d_ops, t->ode.ops.data(), sizeof(tb_op_t) * d.nops, cudaMemcpyHostToDevice));
        TB_CUDA(cudaMalloc(&t->d_deg, nc));
        TB_CUDA(cudaMemcpy(t->d_deg, t->T.deg.data(), nc, cudaMemcpyHostToDevice));
        TB_CUDA(cudaMalloc(&t->d_off, sizeof(int) * (nc + 1)));
        TB_CUDA(cudaMemcpy(t->d_off, t->T.pair_off.data(), sizeof(int) * (nc + 1), cudaMemcpyHostToDevice));
        const size_t np = t->T.pair_a.size();
        TB_CUDA(cudaMalloc(&t->d_pa, 2 * np));
        TB_CUDA(cudaMalloc(&t->d_pb, 2 * np));
        TB_CUDA(cudaMemcpy(t->d_pa, t->T.pair_a.data(), 2 * np, cudaMemcpyHostToDevice));
        TB_CUDA(cudaMemcpy(t->d_pb, t->T.pair_b.data(), 2 * np, cudaMemcpyHostToDevice));
        std::vector<unsigned int> pab(np);
        for (size_t q = 0; q < np; q++) pab[q] = (unsigned int)t->T.pair_a[q] | ((unsigned int)t->T.pair_b[q] << 16);
        TB_CUDA(cudaMalloc(&t->d_pab, 4 * (np ? np : 1)));
        TB_CUDA(cudaMemcpy(t->d_pab, pab.data(), 4 * np, cudaMemcpyHostToDevice));
        TB_CUDA(cudaMalloc(&t->d_sched, 4 * (t->sched.size() ? t->sched.size() : 1)));
        TB_CUDA(cudaMemcpy(t->d_sched, t->sched.data(), 4 * t->sched.size(), cudaMemcpyHostToDevice));
    }
    const int threads = ((nc + 31) / 32) * 32;
    size_t smem = sizeof(double) * nc * 9; /* scratch + 8 staging buffers */
    const size_t npairs = t->T.pair_a.size();
    const bool pairs_in_smem = smem + 4 * npairs <= 100 * 1024;
    if (pairs_in_smem) smem += 4 * npairs;
    const bool fast = t->mode == TB_MODE_FAST;
    if (fast) TB_CUDA(cudaFuncSetAttribute(tn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else TB_CUDA(cudaFuncSetAttribute(tn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int bps = 1, nsm = 148;
    if (fast) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, tn_kernel<true>, threads, smem));
    else TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, tn_kernel<false>, threads, smem));
    TB_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    if (bps < 1) bps = 1;
    if (bps > 16) bps = 16;
    long long blocks = (long long)nsm * bps;
    if (blocks > N) blocks = N;
    TnArgs a{};
    a.N = N; a.dof = d.dof; a.npar = d.npar; a.ntmp = d.ntmp; a.nops = d.nops; a.P = P; a.nc = nc; a.d = t->varorder;
    a.nser = 2 * d.dof + d.ntmp + 1;
    a.ops = t->d_ops; a.deg = t->d_deg; a.pair_off = t->d_off; a.pair_a = t->d_pa; a.pair_b = t->d_pb; a.pair_ab = t->d_pab; a.sched = t->d_sched; a.nsched = (int)t->sched.size(); a.npairs = (int)npairs; a.pairs_in_smem = pairs_in_smem ? 1 : 0;
    a.nt = trange ? nt : 0; a.t0 = t0; a.tmax = tmax; a.abstol = o->abstol; a.reltol = o->reltol;
    a.minstep = o->minstepsize; a.maxstep = o->maxstepsize; a.maxsteps = o->maxsteps; a.params_per_traj = o->params_per_traj;
    const size_t nq = (size_t)d.dof * nc * N, npar = (size_t)(d.npar > 0 ? d.npar : 1) * (o->params_per_traj ? N : 1);
    double *dq = nullptr, *dp = nullptr, *dtr = nullptr, *dwork = nullptr, *dxf = nullptr, *dtf = nullptr, *dxv = nullptr;
    long long *dns = nullptr; int *dst = nullptr; unsigned long long *dctr = nullptr;
    int rc = TB_OK;
    auto fail = [&](cudaError_t e, const char *what) { if (e != cudaSuccess && rc == TB_OK) { tb::set_error(std::string(what) + ": " + cudaGetErrorString(e)); rc = TB_ERR_CUDA; } };
    fail(cudaMalloc(&dq, 8 * nq), "cudaMalloc q0");
    fail(cudaMalloc(&dp, 8 * npar), "cudaMalloc params");
    fail(cudaMalloc(&dwork, 8 * (size_t)blocks * a.nser * (P + 1) * nc), "cudaMalloc workspace");
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
        if (d.npar > 0 && params) fail(cudaMemcpy(dp, params, 8 * npar, cudaMemcpyHostToDevice), "copy params");
        if (trange) fail(cudaMemcpy(dtr, trange, 8 * nt, cudaMemcpyHostToDevice), "copy trange");
        fail(cudaMemset(dctr, 0, 8), "memset counter");
        fail(cudaMemset(dwork, 0, 8 * (size_t)blocks * a.nser * (P + 1) * nc), "memset workspace");
        a.q0 = dq; a.params = dp; a.trange = trange ? dtr : nullptr; a.work = dwork; a.xfinal = dxf; a.tfinal = dtf;
        a.nsteps = dns; a.status = dst; a.xv_trange = dxv; a.counter = dctr;
    }
    if (rc == TB_OK) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, 0);
        if (fast) tn_kernel<true><<<(unsigned)blocks, threads, smem>>>(a);
        else tn_kernel<false><<<(unsigned)blocks, threads, smem>>>(a);
        cudaEventRecord(e1, 0);
        tb::count_launch(1);
        fail(cudaGetLastError(), "tn_kernel launch");
        fail(cudaDeviceSynchronize(), "tn_kernel");
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
    cudaFree(dq); cudaFree(dp); cudaFree(dtr); cudaFree(dwork); cudaFree(dxf); cudaFree(dtf); cudaFree(dxv);
    cudaFree(dns); cudaFree(dst); cudaFree(dctr);
    return rc;
}

} /* extern "C" */

/* ---- no-allocation entry point ------------------------------------------------------------------------------
 * tb_tn_integrate_static: the same tn_kernel, for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Everything tb_tn_integrate allocates per call -- the work-queue
 * counter, the device copies of the tape and the index tables, and the jet workspace [blocks][nser][P+1][nc] -- lives in
 * ONE file-scope device array of TB_TN_STATIC_SCRATCH_BYTES; inputs and outputs are the caller's DEVICE buffers.
 * In host code `tn_static_scratch` is only a stand-in for the symbol: its device address is resolved once per device
 * with cudaGetSymbolAddress (that device current) and that pointer is what the kernel receives.
 * Every call shares the array and finds the previous call's contents in it: the counter and the part of the workspace
 * this call uses are cleared on the caller's stream before the launch, the tables are sent again whenever another handle used the array in between.  Two calls
 * that overlap on the GPU (different streams, different host threads) race on it: serialise them.                 */
static __device__ __align__(256) unsigned char tn_static_scratch[TB_TN_STATIC_SCRATCH_BYTES];

namespace {
std::mutex tn_static_mu;
std::map<int, unsigned char *> tn_static_addr; /* device -> device address of tn_static_scratch */
std::map<int, unsigned long long> tn_static_owner; /* device -> serial of the handle whose tables are in the array */

int tn_static_base(int dev, unsigned char **base) {
    std::lock_guard<std::mutex> lk(tn_static_mu);
    auto it = tn_static_addr.find(dev);
    if (it == tn_static_addr.end()) {
        unsigned char *ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, tn_static_scratch); /* device `dev` is current */
        if (e != cudaSuccess || !ptr) {
            tb::set_error(std::string("tb_tn_integrate_static: cudaGetSymbolAddress(tn_static_scratch): ") + cudaGetErrorString(e));
            return TB_ERR_CUDA;
        }
        it = tn_static_addr.emplace(dev, ptr).first;
    }
    *base = it->second;
    return TB_OK;
}
} // namespace

extern "C" int tb_tn_integrate_static(tb_tn_t *t, int64_t N, const double *d_q0, const double *d_params, double t0, double tmax,
                                      const double *d_trange, int64_t nt, const tb_opts_t *o, const tb_tn_out_t *d_out,
                                      void *stream) {
    if (!t || !d_q0 || !o || !d_out || N < 0 || (d_trange && nt < 2)) { tb::set_error("tb_tn_integrate_static: bad argument"); return TB_ERR_ARG; }
    const tb_ode_desc_t &d = t->ode.desc;
    if (d.npar > 0 && !d_params) { tb::set_error("tb_tn_integrate_static: the ODE has parameters, d_params is NULL"); return TB_ERR_ARG; }
    if (N == 0) return TB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, nsm = 148, bps = 1;
    TB_CUDA(cudaGetDevice(&dev));
    const int nc = t->T.ncoef, P = t->order;
    const size_t np = t->T.pair_a.size();
    const int threads = ((nc + 31) / 32) * 32;
    size_t smem = sizeof(double) * nc * 9;
    const bool pairs_in_smem = smem + 4 * np <= 100 * 1024;
    if (pairs_in_smem) smem += 4 * np;
    const bool fast = t->mode == TB_MODE_FAST;
    if (fast) TB_CUDA(cudaFuncSetAttribute(tn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else TB_CUDA(cudaFuncSetAttribute(tn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (fast) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, tn_kernel<true>, threads, smem));
    else TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, tn_kernel<false>, threads, smem));
    TB_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    bps = bps < 1 ? 1 : (bps > 16 ? 16 : bps);
    /* layout of the array: counter | tape | degrees | pair offsets | packed pairs | pair_a | pair_b | workspace */
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const int nser = 2 * d.dof + d.ntmp + 1;
    const size_t o_ctr = 0, o_ops = 256, o_deg = o_ops + up(sizeof(tb_op_t) * (size_t)d.nops), o_off = o_deg + up((size_t)nc),
                 o_pab = o_off + up(sizeof(int) * (size_t)(nc + 1)), o_pa = o_pab + up(4 * np), o_pb = o_pa + up(2 * np),
                 o_sch = o_pb + up(2 * np), o_work = o_sch + up(4 * t->sched.size());
    const size_t per_block = sizeof(double) * (size_t)nser * (size_t)(P + 1) * (size_t)nc;
    if (o_work + per_block > (size_t)TB_TN_STATIC_SCRATCH_BYTES) {
        tb::set_error("tb_tn_integrate_static: one trajectory needs " + std::to_string(o_work + per_block) +
                      " bytes of scratch (tables + " + std::to_string(nser) + " jets x " + std::to_string(P + 1) + " x " +
                      std::to_string(nc) + " coefficients), more than TB_TN_STATIC_SCRATCH_BYTES = " +
                      std::to_string((size_t)TB_TN_STATIC_SCRATCH_BYTES) + "; use tb_tn_integrate");
        return TB_ERR_UNSUPPORTED;
    }
    long long blocks = (long long)nsm * bps;
    if (blocks > N) blocks = N;
    const long long fit = (long long)(((size_t)TB_TN_STATIC_SCRATCH_BYTES - o_work) / per_block);
    if (blocks > fit) blocks = fit; /* fewer resident CTAs than tb_tn_integrate would use: same results, the queue is shared */
    unsigned char *base = nullptr;
    const int src = tn_static_base(dev, &base);
    if (src != TB_OK) return src;
    TB_CUDA(cudaMemsetAsync(base + o_ctr, 0, 256, st));
    TB_CUDA(cudaMemsetAsync(base + o_work, 0, per_block * (size_t)blocks, st));
    /* the tape and the index tables are sent when another handle (or nobody) used the array last on this device: the
     * calls after the first of a handle enqueue two clears and the kernel, nothing else */
    bool resident;
    {
        std::lock_guard<std::mutex> lk(tn_static_mu);
        resident = tn_static_owner.count(dev) && tn_static_owner[dev] == t->serial;
        tn_static_owner[dev] = t->serial;
    }
    if (!resident) {
        /* packed pairs (host-side scratch of this thread; pageable copies are staged before cudaMemcpyAsync returns) */
        static thread_local std::vector<unsigned int> pab;
        pab.resize(np);
        for (size_t q = 0; q < np; q++) pab[q] = (unsigned int)t->T.pair_a[q] | ((unsigned int)t->T.pair_b[q] << 16);
        if (d.nops > 0) TB_CUDA(cudaMemcpyAsync(base + o_ops, t->ode.ops.data(), sizeof(tb_op_t) * (size_t)d.nops, cudaMemcpyHostToDevice, st));
        TB_CUDA(cudaMemcpyAsync(base + o_deg, t->T.deg.data(), (size_t)nc, cudaMemcpyHostToDevice, st));
        TB_CUDA(cudaMemcpyAsync(base + o_off, t->T.pair_off.data(), sizeof(int) * (size_t)(nc + 1), cudaMemcpyHostToDevice, st));
        if (np > 0) {
            TB_CUDA(cudaMemcpyAsync(base + o_pab, pab.data(), 4 * np, cudaMemcpyHostToDevice, st));
            TB_CUDA(cudaMemcpyAsync(base + o_pa, t->T.pair_a.data(), 2 * np, cudaMemcpyHostToDevice, st));
            TB_CUDA(cudaMemcpyAsync(base + o_pb, t->T.pair_b.data(), 2 * np, cudaMemcpyHostToDevice, st));
        }
        if (!t->sched.empty()) TB_CUDA(cudaMemcpyAsync(base + o_sch, t->sched.data(), 4 * t->sched.size(), cudaMemcpyHostToDevice, st));
    }
    TnArgs a{};
    a.N = N; a.dof = d.dof; a.npar = d.npar; a.ntmp = d.ntmp; a.nops = d.nops; a.P = P; a.nc = nc; a.d = t->varorder; a.nser = nser;
    a.ops = (const tb_op_t *)(base + o_ops); a.deg = base + o_deg; a.pair_off = (const int *)(base + o_off);
    a.pair_a = (const unsigned short *)(base + o_pa); a.pair_b = (const unsigned short *)(base + o_pb);
    a.pair_ab = (const unsigned int *)(base + o_pab); a.npairs = (int)np;
    a.sched = (const unsigned int *)(base + o_sch); a.nsched = (int)t->sched.size(); a.pairs_in_smem = pairs_in_smem ? 1 : 0;
    a.q0 = d_q0; a.params = d_params; a.trange = d_trange; a.nt = d_trange ? nt : 0; a.t0 = t0; a.tmax = tmax;
    a.abstol = o->abstol; a.reltol = o->reltol; a.minstep = o->minstepsize; a.maxstep = o->maxstepsize; a.maxsteps = o->maxsteps;
    a.params_per_traj = o->params_per_traj;
    a.work = (double *)(base + o_work);
    a.xfinal = d_out->xfinal; a.tfinal = d_out->tfinal; a.nsteps = (long long *)d_out->nsteps; a.status = d_out->status;
    a.xv_trange = d_trange ? d_out->xv_trange : nullptr;
    a.counter = (unsigned long long *)(base + o_ctr);
    if (fast) tn_kernel<true><<<(unsigned)blocks, threads, smem, st>>>(a);
    else tn_kernel<false><<<(unsigned)blocks, threads, smem, st>>>(a);
    tb::count_launch(1);
    TB_CUDA(cudaGetLastError());
    return TB_OK; /* asynchronous on `stream`: the caller synchronises before reading the outputs */
}
