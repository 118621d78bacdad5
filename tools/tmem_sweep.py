#!/usr/bin/env python3
"""Kepler-28 parity kernel with the tensor-memory tier: bit-parity against the oracle on a small ensemble, then the
bench workload (2^20 trajectories x 100 periods by default) for several register budgets.  usage: tools/tmem_sweep.py [ntraj]"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
import oracle

import json
ntraj = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
cfg_file = sys.argv[2] if len(sys.argv) > 2 else None
modes = [int(m) for m in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 1]
tape = P.kepler_tape()
lib = L.lib()
dev = torch.device("cuda", 0)
q0, mu, _ = P.kepler_ensemble(ntraj)
tmax = 2 * np.pi * 100
d_q0, d_mu = torch.from_numpy(q0).to(dev), torch.from_numpy(mu).to(dev)
d_xf = torch.empty((4, ntraj), dtype=torch.float64, device=dev)
d_tf = torch.empty(ntraj, dtype=torch.float64, device=dev)
d_ns = torch.empty(ntraj, dtype=torch.int64, device=dev)
d_st = torch.empty(ntraj, dtype=torch.int32, device=dev)
d_work = torch.zeros(256, dtype=torch.uint8, device=dev)
o = L.TbOut()
o.xfinal = C.cast(d_xf.data_ptr(), L._DP); o.tfinal = C.cast(d_tf.data_ptr(), L._DP)
o.nsteps = C.cast(d_ns.data_ptr(), C.POINTER(C.c_int64)); o.status = C.cast(d_st.data_ptr(), C.POINTER(C.c_int32))
opts = tb.make_opts(abstol=1e-20, maxsteps=10000)
qs, _, _ = P.kepler_ensemble(300)
ref = oracle.integrate(tape, 28, qs, 0.0, 2 * np.pi * 3, mu, maxsteps=2000, nthreads=8)
base_x = None
configs = [("shipped (4 reg, smem)", dict(flags=262144)),
           ("tmem, 4 reg", dict(flags=262144 | 524288, reg_budget=4 * 29)),
           ("tmem, 3 reg", dict(flags=262144 | 524288, reg_budget=3 * 29)),
           ("tmem, 2 reg", dict(flags=262144 | 524288, reg_budget=2 * 29)),
           ("tmem, 3 reg, 2 tmem series", dict(flags=262144 | 524288, reg_budget=3 * 29, tmem_budget=2 * 29)),
           ("tmem, 3 reg, 1 tmem series", dict(flags=262144 | 524288, reg_budget=3 * 29, tmem_budget=29))]
if cfg_file:
    configs = [(c["name"], c["kw"]) for c in json.load(open(cfg_file))]
for mode in modes:
    for name, kw in configs:
        kw = dict(kw)
        opts.sort_by_cost = int(kw.pop("sort_by_cost", 0))
        try:
            k = tb.Kernel(tape, 28, L.KIND_PLAIN, mode, **kw)
        except Exception as e:
            print(name, "compile failed:", str(e)[:200]); continue
        info = k.info()
        ok = None
        if mode == L.MODE_PARITY:
            g = tb.taylorinteg(tape, qs, 0.0, 2 * np.pi * 3, 28, 1e-20, mu, maxsteps=2000, kernel_opts=dict(kw))
            ok = bool(np.array_equal(g.xfinal, ref.xfinal) and np.array_equal(g.nsteps, ref.nsteps))
        ms = []
        for it in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            L.check(lib.tb_integrate_device(k.handle, ntraj, d_q0.data_ptr(), d_mu.data_ptr(), 0.0, tmax, None, 0, C.byref(opts),
                                            C.byref(o), d_work.data_ptr(), torch.cuda.current_stream().cuda_stream), "dev")
            e1.record(); torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        steps = int(d_ns.sum().item())
        x = d_xf.cpu().numpy()
        if base_x is None:
            base_x = x.copy()
        same = bool(np.array_equal(x, base_x)) if mode == L.MODE_PARITY else None
        print(f"mode {mode} {name:32s} regs {info['regs']} local {info['local_bytes']} smem {info['smem_bytes']} "
              f"parity_small {ok} same_as_shipped {same} ms {min(ms):.1f} steps/s {steps / min(ms) * 1e3:.4g} all_ok {bool((d_st == 0).all())}", flush=True)
