/* tb_internal.h -- shared declarations of libtaylorb200's host side */
#ifndef TB_INTERNAL_H
#define TB_INTERNAL_H

#include <map>
#include <string>
#include <vector>

#include "../../include/taylorb200.h"

namespace tb {

/* validated, owned copy of a tape description */
struct Ode {
    tb_ode_desc_t desc;
    std::vector<tb_op_t> ops;
    std::vector<tb_jac_entry_t> jac;
};

/* result of the storage analysis of tb_codegen.cpp */
struct Layout {
    int nx = 0;                   /* state jets (dof, or dof + dof^2 for Lyapunov) */
    int nslot = 0;                /* series held in shared memory per trajectory   */
    std::map<int32_t, int> slot;  /* operand -> slot (state jets: slot = index)    */
    int block_threads = 0;
    int blocks_per_sm = 1;
    size_t smem_bytes = 0;        /* dynamic shared memory per block               */
    int variant = 0;              /* which generator produced the kernel           */
};

int generate(const Ode &ode, const tb_compile_opts_t &co, Layout &L, std::string &src, std::string &err);
/* variant 2: fully unrolled, register-tiled (tb_codegen_unrolled.cpp); TB_ERR_UNSUPPORTED when it does not apply */
int generate_unrolled(const Ode &ode, const tb_compile_opts_t &co, Layout &L, std::string &src, std::string &err);
/* picks the generator: variant 2 where it applies unless TB_CF_ROLLED is set */
int generate_any(const Ode &ode, const tb_compile_opts_t &co, Layout &L, std::string &src, std::string &err);

/* device source text embedded at build time (tools/embed.py -> tb_embedded.inc) */
const char *embedded_math_h();
const char *embedded_device_pre();
const char *embedded_device_drv();

void set_error(const std::string &msg);
/* NVRTC -> sm_100a cubin with the in-tree disk cache (tb_lib.cpp); module / launch helpers over the driver API */
int compile_cubin(const std::string &source, int mode, bool disk_cache, std::vector<char> &cubin, std::string &log);
int module_load(const std::vector<char> &cubin, const char *fname, size_t smem, void **mod, void **fn);
void module_unload(void *mod);
int launch_fn(void *fn, unsigned grid, unsigned block, size_t smem, void *stream, void *arg_struct);
const char *embedded_wide2_dev();
void set_last_kernel_ms(double ms);

} // namespace tb

#endif
