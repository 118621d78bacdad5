/*
 * taylorb200.h -- C ABI of libtaylorb200.so
 *
 * B200-native (sm_100a) replacement for the ensemble hot path of
 * TaylorIntegration.jl: the `taylorinteg` / `taylorstep!` driver loop run over
 * N independent trajectories.  Plain pointers and sizes only; no torch types.
 *
 * The reference has no FFI: its seam is Julia multiple dispatch.  Every entry
 * point below names the reference method(s) it stands in for (paths relative
 * to the TaylorIntegration.jl source tree, v0.18.14):
 *
 *   tb_ode_create        <- what `@taylorize` emits: the body of
 *                           `jetcoeffs!(::Val{f}, t, x, dx, params, rv)`
 *                           src/parse_eqs.jl:236-358 (skeleton 259-264),
 *                           one TaylorSeries.<op>!(res, a, b, ord) call per
 *                           tape entry (parse_eqs.jl:963-1135)
 *   tb_compile           <- `_determine_parsing!` + `_allocate_jetcoeffs!`
 *                           src/integrator/jetcoeffs.jl:147-192,
 *                           src/parse_eqs.jl:138-181
 *   tb_integrate*        <- `taylorinteg!(dense, f!, q0, t0, tmax, abstol,
 *                           cache, params; ...)`
 *                           src/integrator/taylorinteg.jl:329-381   (t0,tmax)
 *                           src/integrator/taylorinteg.jl:582-642   (trange)
 *                           src/rootfinding.jl:347-438              (events)
 *                           src/lyapunovspectrum.jl:340-429         (Lyapunov)
 *                           batched over the trajectory index.
 *   tb_jet               <- one `taylorstep!` src/integrator/taylorstep.jl:59-82
 *                           (jets + step size of the first step; parity probe)
 *   tb_dense_eval        <- `(sol::TaylorSolution)(t)`
 *                           src/integrator/taylorsolution.jl:337-423
 *
 * Layout convention: every per-trajectory array is structure-of-arrays with
 * the trajectory index fastest, i.e. element (component c, trajectory i) of a
 * `[dof x N]` array lives at `c*N + i`.
 *
 * Error convention: functions return 0 on success and a negative tb_status on
 * API / compile / launch failure (message through tb_last_error()).
 * Integration-level conditions never fail the call: they are reported per
 * trajectory in `status[i]` as TB_ST_* bit flags mirroring the reference's
 * warnings.
 */
#ifndef TAYLORB200_H
#define TAYLORB200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB_ABI_VERSION 2

/* ---- return codes -------------------------------------------------------- */
enum tb_status {
    TB_OK = 0,
    TB_ERR_ARG = -1,      /* bad argument / malformed tape                   */
    TB_ERR_COMPILE = -2,  /* NVRTC failed (log in tb_last_error)             */
    TB_ERR_CUDA = -3,     /* CUDA runtime error or no usable device          */
    TB_ERR_UNSUPPORTED = -4
};

/* ---- per-trajectory status bits ------------------------------------------ */
#define TB_ST_OK 0
#define TB_ST_MAXSTEPS 1   /* "Maximum number of integration steps reached"  taylorinteg.jl:372-377 */
#define TB_ST_ZERO_STEP 2  /* "The step-size is zero; aborting integration." taylorinteg.jl:360-363 */
#define TB_ST_NR_NOCONV 4  /* "Newton-Raphson did not converge ..."          rootfinding.jl:222-224 */
#define TB_ST_NONFINITE 8  /* state became NaN/Inf (no reference analogue; the reference would loop) */
#define TB_ST_EVENT_OVERFLOW 16 /* more events than `event_capacity`        */

/* ---- tape ----------------------------------------------------------------
 * One entry = one `TaylorSeries.<op>!(res, a, b, ord)` call in the body of the
 * generated `for ord = 0:order-1` loop (parse_eqs.jl:259-264).
 */
enum tb_opcode {
    TB_OP_NOP = 0,
    TB_OP_IDENTITY = 1, /* identity!(res, a, k)                 res[k] = a[k]                 */
    TB_OP_NEG = 2,      /* subst!(res, a, k)                    res[k] = -a[k]                */
    TB_OP_ADD = 3,      /* add!(res, a, b, k)     (series|scalar) + (series|scalar)           */
    TB_OP_SUB = 4,      /* subst!(res, a, b, k)                                                */
    TB_OP_MUL = 5,      /* mul!(res, a, b, k)     Cauchy product, or scalar*series            */
    TB_OP_DIV = 6,      /* div!(res, a, b, k)     series/series, series/scalar, scalar/series */
    TB_OP_SQR = 7,      /* sqr!(res, a, k)        (what pow! dispatches to for ^2)            */
    TB_OP_SQRT = 8,     /* sqrt!(res, a, k)                                                    */
    TB_OP_POW = 9,      /* pow!(res, a, aux, imm, k)  constant real exponent in .imm          */
    TB_OP_EXP = 10,     /* exp!(res, a, k)                                                     */
    TB_OP_LOG = 11,     /* log!(res, a, k)                                                     */
    TB_OP_SINCOS = 12,  /* sincos!(res, aux, a, k)    res = sin(a), aux = cos(a)              */
    TB_OP_ZERO = 13,    /* zero!(res, a, k)                                                    */
    TB_OP_ONE = 14,     /* one!(res, a, k)                                                     */
    TB_OP_TAN = 15,     /* tan!(res, a, aux, k)       aux = tan(a)^2                          */
    TB_OP_SINHCOSH = 16,/* sinhcosh!(res, aux, a, k)  res = sinh(a), aux = cosh(a)            */
    TB_OP_TANH = 17,    /* tanh!(res, a, aux, k)      aux = tanh(a)^2                         */
    TB_OP_ATAN = 18,    /* atan!(res, a, aux, k)      aux = 1 + a^2                           */
    TB_OP_ASIN = 19,    /* asin!(res, a, aux, k)      aux = sqrt(1 - a^2)                     */
    TB_OP_ACOS = 20,    /* acos!(res, a, aux, k)      aux = sqrt(1 - a^2)                     */
    TB_OP__COUNT
};

/* operand = kind<<24 | index */
enum tb_operand_kind {
    TB_K_NONE = 0,
    TB_K_X = 1,    /* x[index]   state jet                                   */
    TB_K_DX = 2,   /* dx[index]  RHS jet                                     */
    TB_K_TMP = 3,  /* temporary series (RetAlloc.v0, parse_eqs.jl:80-107)    */
    TB_K_PAR = 4,  /* params[index], a plain number (`local mu = p`)         */
    TB_K_TIME = 5, /* the independent variable t = [t0, 1, 0, ...]           */
    TB_K_IMM = 6   /* numeric literal, value in tb_op_t.imm                  */
};
#define TB_OPERAND(kind, index) ((int32_t)(((uint32_t)(kind) << 24) | ((uint32_t)(index) & 0xFFFFFFu)))
#define TB_OPERAND_KIND(o) ((int)(((uint32_t)(o)) >> 24))
#define TB_OPERAND_INDEX(o) ((int)(((uint32_t)(o)) & 0xFFFFFFu))

typedef struct tb_op {
    int32_t op;   /* tb_opcode                                               */
    int32_t res;  /* result operand (TB_K_DX or TB_K_TMP)                    */
    int32_t a;    /* first argument                                          */
    int32_t b;    /* second argument or TB_K_NONE                            */
    int32_t aux;  /* auxiliary series (pow!, sincos!, ...) or TB_K_NONE      */
    int32_t flags;
    double imm;   /* literal operand or pow! exponent                        */
} tb_op_t;

/* Jacobian entry for Lyapunov spectra (stabilitymatrix!, lyapunovspectrum.jl:45-80).
 * The order-1 TaylorN pass over the tape is done at tape-build time; each entry
 * of the dof x dof Jacobian (row-major here: jac[i*dof + j] = d(dx_i)/d(x_j))
 * is structurally zero, a constant series [c,0,0,...] or a series operand.    */
enum tb_jac_kind {
    TB_JAC_ZERO = 0,
    TB_JAC_CONST = 1,  /* the constant series [c, 0, 0, ...]                  */
    TB_JAC_SERIES = 2, /* a series operand                                    */
    TB_JAC_PAR = 3     /* the constant series [c * params[index], 0, ...]; operand = TB_OPERAND(TB_K_PAR, index) */
};
typedef struct tb_jac_entry {
    int32_t kind;    /* tb_jac_kind                                          */
    int32_t operand; /* series operand (TB_JAC_SERIES) or parameter (TB_JAC_PAR) */
    double c;        /* constant (TB_JAC_CONST) or multiplier (TB_JAC_PAR)   */
} tb_jac_entry_t;

/* Event gate: the Bool of the `(Bool, Taylor1)` tuple returned by g
 * (rootfinding.jl:124-132), restricted to a sign test on a constant term.    */
enum tb_gate_cmp { TB_GATE_ALWAYS = 0, TB_GATE_GT0 = 1, TB_GATE_LT0 = 2, TB_GATE_GE0 = 3, TB_GATE_LE0 = 4 };

typedef struct tb_ode_desc {
    int32_t dof;      /* number of state components                          */
    int32_t npar;     /* number of scalar parameters                         */
    int32_t ntmp;     /* number of temporaries (all three tape segments)     */
    int32_t nops;     /* RHS tape: ops[0 .. nops)                            */
    int32_t ev_nops;  /* event tape: ops[nops .. nops+ev_nops)   (0 = none)  */
    int32_t jac_nops; /* Jacobian tape: the next jac_nops ops    (0 = none)  */
    const tb_op_t *ops;
    int32_t ev_g;        /* operand holding the series of g                  */
    int32_t ev_gate;     /* operand whose constant term gates the event      */
    int32_t ev_gate_cmp; /* tb_gate_cmp                                      */
    int32_t has_jac;     /* jac[] valid (dof*dof entries)                    */
    const tb_jac_entry_t *jac;
} tb_ode_desc_t;

/* ---- compile options ----------------------------------------------------- */
enum tb_mode {
    TB_MODE_PARITY = 0, /* reference operation order, no FMA contraction,
                           correctly rounded div/sqrt, shared deterministic
                           pow/log/exp/sincos: bitwise equal to the oracle    */
    TB_MODE_FAST = 1    /* FMA, re-associated sums, reciprocal multiplies     */
};
enum tb_kind {
    TB_KIND_PLAIN = 0,  /* taylorinteg (t0,tmax) and trange                   */
    TB_KIND_EVENTS = 1, /* + findroot!                                        */
    TB_KIND_LYAP = 2,   /* lyap_taylorinteg                                   */
    TB_KIND_LYAPCOL = 3 /* lyap_taylorinteg, column-parallel: the ODE handed to tb_compile is the COLUMN SYSTEM
                         * [x; v] of 2*n components (x' = f(x), v' = J(x) v, built by the front end from the tape
                         * and its Jacobian); a team of n lanes integrates one trajectory, lane j carrying column j
                         * of the fundamental matrix, and the per-step modified Gram-Schmidt runs over warp
                         * shuffles.  q0, xfinal, lambda are [n x N] as for TB_KIND_LYAP; same bits.            */
};

typedef struct tb_compile_opts {
    int32_t order;         /* Taylor order p (1..40)                          */
    int32_t mode;          /* tb_mode                                         */
    int32_t kind;          /* tb_kind                                         */
    int32_t block_threads; /* 0 = default                                     */
    int32_t reg_budget;    /* variant 2: doubles of jet storage kept in registers per trajectory (0 = default: 3 series, <0 = none) */
    int32_t min_blocks_per_sm; /* 0 = default                                 */
    int32_t flags;         /* TB_CF_*                                         */
    int32_t tmem_budget;   /* with TB_CF_TMEM: doubles of jet storage in tensor memory per trajectory (0 = as many as fit) */
} tb_compile_opts_t;
#define TB_CF_NO_DISK_CACHE 1 /* do not read/write the on-disk cubin cache    */
#define TB_CF_TMEM_TAILS 2    /* parity TB_KIND_PLAIN + TB_CF_FINAL_ONLY: state jets that only Horner / the step size read live in tensor memory */
#define TB_CF_ROLLED 4        /* force kernel variant 1 (rolled loops, all jets in shared memory) */
/* tuning switches of the unrolled generator (variant 2); all default off */
#define TB_CF_POW_RECIP 8      /* parity: pow! finishes with reciprocal + Markstein correction instead of a full division (same bits; measured -2.5 % on
                                  Kepler-28: three more values live across the group spill).  Bit of the retired TB_CF_NO_MEMBAR. */
#define TB_CF_NO_PAIRING 16   /* do not emit two consecutive orders of a recurrence together */
#define TB_CF_NO_GROUPING 32  /* one recurrence entry per group (no interleaving of independent sums) */
#define TB_CF_NO_QUICK 64     /* parity mode: single jet body with guarded divisions */
#define TB_CF_STORE_X 256     /* fast mode: keep state jets stored even when they can be derived from dx */
#define TB_CF_DERIVE_X 512    /* parity mode: derive unread state jets from dx (default: stored) */
#define TB_CF_BLOCK_LOCKSTEP 128 /* barrier per order: the warps of a block share one instruction stream */
#define TB_CF_STEP_LOCKSTEP 1024 /* one block barrier per step (no barriers inside the jet body) */
#define TB_CF_NO_TRIM_X 4096    /* parity mode: store x[0..2] of tail-only state jets as well */
#define TB_CF_SPLIT 8192        /* registers hold the low coefficients of every recurrence-read series, shared memory the high ones */
#define TB_CF_NO_STEP_LOCKSTEP 16384 /* event kernels: no block barrier per step (it is their default) */
#define TB_CF_PAIR_POW 65536     /* parity: evaluate the two step-size powers in one interleaved body (measured slower) */
#define TB_CF_LYAP_TRANGE 131072  /* TB_KIND_LYAPCOL: build the trange variant (tb_integrate_trange needs it) */
#define TB_CF_FINAL_ONLY 262144   /* TB_KIND_PLAIN: final-state outputs only (no trange, tv/xv, psol) */
#define TB_CF_TMEM 524288        /* TB_KIND_PLAIN: tensor memory as a third storage tier for recurrence-read series (tcgen05.ld/st) */
#define TB_CF_TOPTRIM(t) (((t) & 15) << 20) /* register series: their top t coefficients live in shared memory */
#define TB_CF_PREFETCH(d) (((d) & 15) << 24) /* storage loads are emitted 2*d source lines before their first use */
#define TB_CF_MULAHEAD(d) (((d) & 7) << 28)  /* parity: the product of a sum term is emitted 2*d lines before its add (software pipelining) */
#define TB_CF_NO_SWPIPE 32768     /* parity: no software pipelining of loads / products (default: PREFETCH(8), MULAHEAD(4)) */
#define TB_CF_SYNC_EVERY4 2048   /* with TB_CF_BLOCK_LOCKSTEP: barrier every 4th order only */

/* ---- integration options: keyword arguments of taylorinteg -------------- */
typedef struct tb_opts {
    double abstol;       /* positional `abstol`                               */
    double reltol;       /* reltol = 0               taylorinteg.jl:313-319   */
    double minstepsize;  /* minstepsize = 0                                   */
    double maxstepsize;  /* maxstepsize = Inf                                 */
    int64_t maxsteps;    /* maxsteps = 500                                    */
    int32_t eventorder;  /* eventorder = 0           rootfinding.jl:310-318   */
    int32_t newtoniter;  /* newtoniter = 10                                   */
    double nrabstol;     /* nrabstol = eps(Float64)                           */
    int32_t event_capacity; /* events kept per trajectory (reference: maxsteps+1) */
    int32_t params_per_traj; /* 1: params is [npar x N]; 0: params is [npar], shared */
    int64_t steps_stride;  /* per-step outputs: columns per trajectory (>= maxsteps+1), 0 = maxsteps+1 */
    int32_t sort_by_cost;  /* scheduling hint: 1 = cost grows with the trajectory index, hand out index N-1 first
                            * (a lane that refills late with an expensive trajectory runs alone at the end); results unchanged */
    int32_t reserved;
} tb_opts_t;

/* ---- outputs (caller-owned; NULL = not requested) ------------------------
 * Host pointers for tb_integrate, device pointers for tb_integrate_device.  */
typedef struct tb_out {
    double *xfinal;    /* [dof x N]   state at the end of the integration     */
    double *tfinal;    /* [N]         time reached                            */
    int64_t *nsteps;   /* [N]         steps taken (reference `nsteps - 1`)    */
    int32_t *status;   /* [N]         TB_ST_* flags                           */
    /* trange variant: taylorinteg.jl:582-642, NaN-prefilled cache.jl:313-316 */
    double *xv_trange; /* [dof x nt x N]  element (c, j, i) at (c*nt + j)*N + i */
    /* per-step storage, the reference's tv / xv (taylorinteg.jl:366-371)     */
    double *tv;        /* [stride x N]      tv[s*N + i], s = 0..nsteps        */
    double *xv;        /* [dof x stride x N]                                  */
    double *psol;      /* [dof x (order+1) x stride x N] dense output, element
                          (c, k, s, i) at ((c*(order+1) + k)*stride + s)*N + i */
    /* events: rootfinding.jl:225-231                                         */
    double *tevents;   /* [event_capacity x N]                                */
    double *xevents;   /* [dof x event_capacity x N]                          */
    double *gresids;   /* [event_capacity x N]                                */
    int32_t *nevents;  /* [N]                                                 */
    /* Lyapunov: lyapunovspectrum.jl:411-415                                  */
    double *lambda;    /* [dof x N]   exponents at tfinal                     */
    double *lambda_v;  /* [dof x stride x N] per step (with tv/xv), optional  */
    /* Lyapunov "continue from state" (the state of a run is (t0, x0, λtsum, Q); the reference keeps it
     * in the cache of lyap_taylorinteg!, lyapunovspectrum.jl:340-429, and has no resume call).  (t0, tmax) drivers only.
     * Rows [(dof*dof + dof + 1) x N], element (row, i) at row*N + i: rows c*dof + r = Q[r, c] (the orthonormal frame the
     * variational state was reset to after the last step), rows dof*dof + c = λtsum[c], last row = t00 (start of the
     * time span the exponents are averaged over).  A second call with q0 = xfinal, t0 = tfinal, lyap_state0 = the
     * lyap_state of the first continues the run: same arithmetic as one run that had been asked to stop at tfinal. */
    double *lyap_state;        /* out, optional */
    const double *lyap_state0; /* in, optional: NULL = fresh run (Q = I, λtsum = 0, t00 = t0) */
} tb_out_t;

typedef struct tb_ode tb_ode_t;       /* validated tape                       */
typedef struct tb_kernel tb_kernel_t; /* compiled kernel (immutable, shareable) */

/* ---- entry points -------------------------------------------------------- */
int tb_abi_version(void);
const char *tb_last_error(void);

/* copies the description; validates operand kinds/indices and op arity      */
tb_ode_t *tb_ode_create(const tb_ode_desc_t *desc);
void tb_ode_free(tb_ode_t *ode);

/* default-initialise option structs with the reference's defaults           */
void tb_opts_default(tb_opts_t *o);
void tb_compile_opts_default(tb_compile_opts_t *o);

/* Generates CUDA C++ for the tape (works without a GPU). Returns a pointer
 * owned by the library (valid until the next call on this thread).          */
const char *tb_generate_source(const tb_ode_t *ode, const tb_compile_opts_t *co);

/* NVRTC-compiles for sm_100a (works without a GPU: cubin only); the module
 * is loaded lazily on first launch. `device` < 0 = current device.          */
tb_kernel_t *tb_compile(const tb_ode_t *ode, const tb_compile_opts_t *co);
void tb_kernel_free(tb_kernel_t *k);
const char *tb_kernel_source(const tb_kernel_t *k);
const char *tb_kernel_log(const tb_kernel_t *k);      /* NVRTC / ptxas -v log */
int tb_kernel_info(const tb_kernel_t *k, int *regs, int *smem_bytes, int *local_bytes,
                   int *block_threads, int *blocks_per_sm);
/* size and bytes of the cubin (for offline cuobjdump)                        */
const void *tb_kernel_cubin(const tb_kernel_t *k, size_t *nbytes);

/* (t0,tmax) ensemble, HOST buffers: copies in, launches, copies out.
 * q0 [dof x N]; params [npar x N] or [npar]; t0/tmax scalars.               */
int tb_integrate(tb_kernel_t *k, int64_t N, const double *q0, const double *params,
                 double t0, double tmax, const tb_opts_t *opts, tb_out_t *out);
/* same with a sorted output grid trange[nt] shared by all trajectories (states in out->xv_trange [dof x nt x N],
 * NaN-prefilled like the reference's caches).  Replaces taylorinteg.jl:582-642 (TB_KIND_PLAIN), rootfinding.jl:485-584
 * (TB_KIND_EVENTS: events as in tb_integrate) and lyapunovspectrum.jl:481-596 (TB_KIND_LYAPCOL: out->lambda_v then
 * holds the exponents at the output times, [dof x nt x N], out->lambda the ones at trange[nt-1]).                     */
int tb_integrate_trange(tb_kernel_t *k, int64_t N, const double *q0, const double *params,
                        const double *trange, int64_t nt, const tb_opts_t *opts, tb_out_t *out);

/* ONE ensemble over several GPUs of this process.  The N trajectories are cut into round-robin blocks
 * of `block` (0 = 4096, or less for small ensembles so that every device gets about 8 blocks); device `devices[d]` takes blocks d, d + ndev, ...; one host thread per device runs the
 * host-buffer path above on its shard and writes its results straight into the caller's [rows x N] buffers.
 * trange == NULL: the (t0, tmax) drivers; otherwise tb_integrate_trange semantics (t0 / tmax ignored).  All outputs of
 * tb_out_t are supported.  tb_last_kernel_ms() afterwards reports the slowest device's kernel.  The reference has no
 * multi-device path; this stands for a `Threads.@threads` loop over taylorinteg calls (docs/src/root_finding.md:121-132). */
int tb_integrate_multi(tb_kernel_t *k, const int32_t *devices, int32_t ndev, int64_t block, int64_t N,
                       const double *q0, const double *params, double t0, double tmax, const double *trange,
                       int64_t nt, const tb_opts_t *opts, tb_out_t *out);

/* DEVICE buffers, asynchronous on `stream` (a cudaStream_t; NULL = default).
 * `work` is a device int64 scratch of tb_work_bytes() bytes.                 */
size_t tb_work_bytes(void);
int tb_integrate_device(tb_kernel_t *k, int64_t N, const double *d_q0, const double *d_params,
                        double t0, double tmax, const double *d_trange, int64_t nt,
                        const tb_opts_t *opts, const tb_out_t *d_out, void *d_work, void *stream);
/* launches issued so far by this library (all kernels), for bench accounting */
int64_t tb_launch_count(void);
/* device time (CUDA events, ms) of the fused kernel launched by the last HOST-buffer entry point
 * (tb_integrate*, tb_tn_integrate, tb_wide_integrate) called on this thread; -1 before any call */
double tb_last_kernel_ms(void);

/* First Taylor step of each trajectory: jets x[dof x (order+1) x N] (element
 * (c,k,i) at (c*(order+1)+k)*N + i) and the step size dt[N]. Host buffers.   */
int tb_jet(tb_kernel_t *k, int64_t N, const double *q0, const double *params, double t0,
           const tb_opts_t *opts, double *xjet, double *dt);

/* Dense-output evaluation sol(t): psol as written by tb_integrate (host),
 * tq[nq] query times; xq [dof x nq x N]. Out-of-range queries give NaN.      */
int tb_dense_eval(int64_t N, int32_t dof, int32_t order, int64_t stride, const double *tv,
                  const double *psol, const int64_t *nsteps, const double *tq, int64_t nq, double *xq);

/* the same on DEVICE buffers (psol, tv, nsteps as the ensemble kernel wrote them; nothing is copied) */
int tb_dense_eval_device(int64_t N, int32_t dof, int32_t order, int64_t stride, const double *d_tv, const double *d_psol,
                         const int64_t *d_nsteps, const double *d_tq, int64_t nq, double *d_xq, void *stream);

/* psol in the ensemble layout -> the flat vector of `TaylorSolutionSerialization` (src/integrator/taylorsolution.jl:
 * 513-557): trajectory i occupies d_out[d_offsets[i] ...], element ((c*nsteps[i] + s)*(order+1) + k) = coefficient k of
 * component c in step s (the order `eachindex(sol.p)` walks the steps x dof matrix, coefficients contiguous).
 * max_nsteps = max_i nsteps[i].  Device buffers; one bandwidth-bound transposing kernel.                        */
int tb_serialize_device(int64_t N, int32_t dof, int32_t order, int64_t stride, int64_t max_nsteps, const double *d_psol,
                        const int64_t *d_nsteps, const int64_t *d_offsets, double *d_out, void *stream);

/* ---- TaylorN jet transport ----------------------------------------------------
 * taylorinteg with U = TaylorN{Float64} (docs/src/pendulum.md:92-108, test/jettransport.jl;
 * solcoeff! src/TaylorIntegration.jl:54-65, _norminf src/integrator/stepsize.jl:26-42):
 * every coefficient of every jet is a polynomial in `nvars` variables truncated at total
 * degree `varorder`, C(nvars+varorder, varorder) numbers stored by ascending degree and,
 * inside a degree, in lexicographically descending exponent order (tb_tn_monomials).
 * q0 / xfinal are [dof x ncoef x N], xv_trange [dof x ncoef x nt x N], trajectory index fastest. */
typedef struct tb_tn tb_tn_t;
typedef struct tb_tn_out {
    double *xfinal;    /* [dof x ncoef x N]                                    */
    double *tfinal;    /* [N]                                                  */
    int64_t *nsteps;   /* [N]                                                  */
    int32_t *status;   /* [N]                                                  */
    double *xv_trange; /* [dof x ncoef x nt x N], NaN-prefilled (cache.jl:313-316) */
} tb_tn_out_t;
tb_tn_t *tb_tn_create(const tb_ode_t *ode, int order, int nvars, int varorder, int mode);
void tb_tn_free(tb_tn_t *t);
int tb_tn_ncoef(const tb_tn_t *t);
/* The kernel runs independent tape entries with a time-index sum (products, squares, powers, divisions) as PAIRS:
 * entry of the schedule = j | 0xffff0000 (tape entry j alone) or j | q << 16 (entries j and q together; the entries
 * between them were moved ahead).  Returns the number of schedule entries, writes at most `cap`.  For inspection. */
int tb_tn_schedule(const tb_tn_t *t, uint32_t *sched, int cap);
/* exponents [ncoef x nvars] of the flat coefficient order; returns ncoef        */
int tb_tn_monomials(const tb_tn_t *t, int32_t *expo);
/* HOST buffers. trange == NULL: (t0,tmax) driver (taylorinteg.jl:329-381); else the
 * trange driver (582-642) on the sorted grid trange[nt].                        */
int tb_tn_integrate(tb_tn_t *t, int64_t N, const double *q0, const double *params, double t0, double tmax,
                    const double *trange, int64_t nt, const tb_opts_t *opts, const tb_tn_out_t *out);

/* The same TaylorN integration for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * q0 / params / trange and every pointer of `d_out` are DEVICE buffers of the caller (layouts as above; xv_trange is
 * used only with a trange); everything else the kernel needs -- work-queue counter, device copies of the tape and the
 * index tables, the jet workspace -- lives in one static device array of TB_TN_STATIC_SCRATCH_BYTES inside the library
 * (every GPU that loads the library holds it until exit).  The call allocates and frees nothing, enqueues clears,
 * (on the first call of a handle, or after another handle used the array) table copies, and ONE kernel on `stream` (a cudaStream_t, NULL = default stream) and returns without synchronising.
 * The grid shrinks to the CTAs whose workspace fits (same results: trajectories come from a shared queue); a problem
 * whose single trajectory does not fit returns TB_ERR_UNSUPPORTED with the sizes in tb_last_error.
 * All calls share the array: calls that overlap on the GPU (two streams, two host threads) race -- serialise them. */
#define TB_TN_STATIC_SCRATCH_BYTES (16u << 20)
int tb_tn_integrate_static(tb_tn_t *t, int64_t N, const double *d_q0, const double *d_params, double t0, double tmax,
                           const double *d_trange, int64_t nt, const tb_opts_t *opts, const tb_tn_out_t *d_out,
                           void *stream);

/* ---- large-state ODEs (one CTA per trajectory) -----------------------------------
 * Same drivers (taylorinteg.jl:329-381, 582-642) for tapes too large for one thread:
 * hundreds of components, 10^4 entries (spin_odes!, benchmark/manyspin_benchmarks.jl:10-41).
 * Outputs: xfinal, tfinal, nsteps, status, xv_trange of the tb_out_t struct, HOST buffers.     */
typedef struct tb_wide tb_wide_t;
tb_wide_t *tb_wide_create(const tb_ode_t *ode, int order, int mode);
void tb_wide_free(tb_wide_t *w);
int tb_wide_info(const tb_wide_t *w, int *nops, int *nlev, int *nser, int *maxwidth);
/* the shared-memory cluster kernel (jets [order][series] in shared memory, one thread per cluster of same-order
 * dependent entries, one barrier per order): clusters, stored series, scalar slots, CTA size, shared memory per CTA,
 * CTAs per SM, and the number of groups of isomorphic clusters that share ONE copy of their entries (0: every cluster
 * carries its own).  TB_ERR_UNSUPPORTED when the stored series do not fit (the global-workspace kernel runs then). */
int tb_wide_layout(const tb_wide_t *w, int *nclusters, int *npersistent, int *ntransient, int *threads, int *smem_bytes,
                   int *ctas_per_sm, int *ngroups);
/* The cluster kernel is specialised per program by NVRTC when its clusters form groups: the group programs become
 * straight-line code with compile-time entries (no dispatch, unrolled coupling sums, the head of each coupling sum's
 * constants in registers).  generated = 0: the interpreter of libtaylorb200.so runs.  tb_wide_source / tb_wide_log:
 * the generated text and the NVRTC / ptxas log. */
int tb_wide_generated(const tb_wide_t *w, int *generated, int *regs, int *local_bytes, int *lc_in_smem);
const char *tb_wide_source(const tb_wide_t *w);
const char *tb_wide_log(const tb_wide_t *w);
int tb_wide_integrate(tb_wide_t *w, int64_t N, const double *q0, const double *params, double t0, double tmax,
                      const double *trange, int64_t nt, const tb_opts_t *opts, const tb_out_t *out);

/* FP64 roofline probe: register-only DFMA chains on the current device.
 * Returns achieved TFLOP/s (2 flops per DFMA) over >= `seconds` of work.     */
int tb_fp64_peak(double seconds, double *tflops, double *ms);

/* device management helpers for hosts without a CUDA binding                 */
int tb_device_count(void);
int tb_set_device(int device);

#ifdef __cplusplus
}
#endif
#endif /* TAYLORB200_H */
