/* tb_wide_prog.h -- the fused internal program of the large-state kernels (tb_wide.cu builds it from the tape;
 * tb_wide2.cu re-partitions it into per-thread clusters with jets in shared memory). */
#ifndef TB_WIDE_PROG_H
#define TB_WIDE_PROG_H

#include <vector>

namespace tbw {

enum { W_NOP = 0, W_LINCOMB = 100, W_SOLCOEFF = 101, W_MUL2 = 102 /* generated kernels only: two products sharing operand b */ };

/* internal tape entry: operands are compact series ids (>= 0) or -1 for a number */
struct WOp {
    int op, res, a, b, aux;
    int para, parb; /* parameter index of a numeric operand, -1 = immediate */
    double imma, immb;
    double imm;     /* pow! exponent */
    int lc_off, lc_n, lc_stride; /* term j of the chain at lc_off + j*lc_stride (chains of a level interleaved: coalesced) */
};

/* the program in TAPE order with its raw (not interleaved) linear-combination terms */
struct Program {
    int order = 0, mode = 0, dof = 0, npar = 0, nser = 0, tser = -1;
    std::vector<WOp> ops;          /* lc_off indexes lc_coef / lc_ser below, lc_stride = 1 */
    std::vector<double> lc_coef;
    std::vector<int> lc_ser, xser, dxser;
};

struct Wide2; /* tb_wide2.cu */
Wide2 *wide2_build(const Program &p, std::string &why);   /* nullptr: the cluster kernel does not apply (why says so) */
void wide2_free(Wide2 *w);
void wide2_info(const Wide2 *w, int *nclusters, int *npersist, int *ntransient, int *threads, int *smem_bytes, int *ctas_per_sm, int *ngroups);
void wide2_gen_info(const Wide2 *w, int *generated, int *regs, int *local_bytes, int *lc_in_smem);
const char *wide2_gen_source(const Wide2 *w);
const char *wide2_gen_log(const Wide2 *w);
int wide2_integrate(Wide2 *w, long long N, const double *q0, const double *params, double t0, double tmax,
                    const double *trange, long long nt, const tb_opts_t *o, const tb_out_t *out);

} // namespace tbw

#endif
