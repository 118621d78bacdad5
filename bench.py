#!/usr/bin/env python3
"""bench.py -- ensemble throughput of the Taylor-integration hot path on N B200s of one node (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W              # this framework (libtaylorb200.so, sm_100a), config 1
    python bench.py --config {1,2,3,4,5a,5b} ...               # the other BASELINE.json configs, same JSON contract
    python bench.py --impl reference --gpus N --steps K ...    # the CPU restatement of the reference on host cores

One "step" = one pass of the hot path over one batch: the whole ensemble integrated from t0 to tmax in ONE kernel
launch.  Config 1 (the headline, default): Kepler order 28, abstol 1e-20, 2^20 eccentricities per GPU x 100 periods,
final-state output.  `value` = trajectory-steps/s with the inputs already resident in HBM (device time, CUDA events);
`e2e` = the same through the host-buffer C-ABI call (pinned host buffers -> device, kernel, results -> host inside the
timed region).  `parity_sample` = the CPU leg's strided sample of the SAME ensemble compared bit for bit with the GPU
results at the same indices; the run exits non-zero when that fails.
Multi-GPU: one process per GPU (torchrun).  Default "weak": every rank integrates its own ensemble; `--scaling strong`
shards ONE ensemble of --ntraj trajectories over the ranks in round-robin blocks.  No data-path collective; the final
NCCL all-gather of the per-rank results is inside the timed region (device arm and e2e arm).
"""
from __future__ import annotations

import argparse
import ctypes as C
import hashlib
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

T_PEND = 7.019250311844546


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="1", choices=["1", "2", "3", "4", "5a", "5b"])
    ap.add_argument("--ntraj", type=int, default=0, help="trajectories per GPU (0 = the BASELINE.json size of the config)")
    ap.add_argument("--periods", type=float, default=100.0, help="config 1: Kepler periods")
    ap.add_argument("--order", type=int, default=0, help="0 = the config's order")
    ap.add_argument("--mode", default="parity", choices=["parity", "fast"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fast", action="store_true", help="skip the extra fast-mode (FMA) measurement (config 1)")
    ap.add_argument("--queue", default="auto", choices=["auto", "forward", "reverse"],
                    help="order the work queue hands trajectories out in (tb_opts_t.sort_by_cost); auto = reverse for configs 1-3")
    ap.add_argument("--reg-budget", type=int, default=0, help="doubles of jet storage in registers (kernel variant 2)")
    ap.add_argument("--block-threads", type=int, default=0)
    ap.add_argument("--kflags", type=int, default=0, help="extra TB_CF_* compile flags")
    ap.add_argument("--full-outputs-kernel", action="store_true",
                    help="config 1: use the build that also carries trange / per-step / dense output code")
    return ap.parse_args()


# ---- clocks sampler ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons, pw = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                pw.append(float(r[6]))
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "power_w_max": max(pw) if pw else None, "samples": len(sm)}


# ---- the six workloads ----------------------------------------------------------------------------------------------
class Workload:
    """One BASELINE.json config: synthetic inputs, the oracle call of the CPU arm, and how to run it on
    the GPU (device-resident through tb_integrate_device where the kernel has that entry point, host buffers else)."""
    key = ""
    metric = ""
    flops_per_step = 0.0
    default_n = 0
    order = 28
    dof = 0

    def __init__(self, a):
        self.a = a
        self.N = a.ntraj or self.default_n
        self.order = a.order or self.order

    # inputs: dict(q0=[dof(,ncoef),N], params=..., per=0/1)
    def describe(self) -> str:
        raise NotImplementedError

    def oracle_run(self, idx, nthreads):
        """Integrates the trajectories `idx` of the ensemble on the CPU oracle; returns a dict of result arrays."""
        raise NotImplementedError

    def cpu_sample(self, target_seconds, nthreads=None):
        """Times the oracle on an evenly strided sample of the ensemble, sized for ~target_seconds."""
        nthreads = nthreads or os.cpu_count() or 1
        n = min(self.N, max(nthreads, self.min_sample(nthreads)))
        rate, secs, res, idx = 0.0, 0.0, None, None
        for _ in range(3):
            idx = np.unique(np.linspace(0, self.N - 1, n).round().astype(np.int64))
            t0 = time.perf_counter()
            res = self.oracle_run(idx, nthreads)
            secs = time.perf_counter() - t0
            rate = int(res["nsteps"].sum()) / secs
            if secs >= 0.5 * target_seconds or idx.size >= self.N:
                break
            n = int(min(self.N, max(n * 2, n * target_seconds / max(secs, 1e-3) * 0.9)))
        desc = (f"{idx.size} of {self.N} trajectories (evenly strided over the ensemble), full horizon each, "
                f"{nthreads} threads, {secs:.1f} s")
        return rate, nthreads, desc, idx, res

    def min_sample(self, nthreads):
        return 4 * nthreads


def _oracle_pack(sol, extra=()):
    d = {"xfinal": sol.xfinal, "nsteps": sol.nsteps, "tfinal": sol.tfinal, "status": sol.status}
    for f in extra:
        d[f] = getattr(sol, f)
    return d


class Kepler(Workload):
    key, metric, flops_per_step, default_n, order, dof = "1", "kepler_order28_ensemble_trajectory_steps_per_s", 4733, 1 << 20, 28, 4

    def __init__(self, a):
        super().__init__(a)
        from taylorintegration_b200 import problems as P
        self.tape = P.kepler_tape()
        self.q0, self.mu, _ = P.kepler_ensemble(self.N)
        self.t0, self.tmax, self.maxsteps = 0.0, 2 * np.pi * a.periods, 10_000
        self.kind, self.params, self.per = 0, self.mu, 0

    def describe(self):
        return (f"kepler1! planar 2-body, order {self.order}, abstol 1e-20, e_i = 0.9*i/(N-1), {self.a.periods:g} periods, "
                f"N = {self.N} trajectories per GPU, output = final state + nsteps")

    def oracle_run(self, idx, nthreads):
        import oracle
        s = oracle.integrate(self.tape, self.order, np.ascontiguousarray(self.q0[:, idx]), self.t0, self.tmax, self.mu,
                             abstol=1e-20, maxsteps=self.maxsteps, nthreads=nthreads)
        return _oracle_pack(s)


class Pendulum(Workload):
    key, metric, flops_per_step, default_n, order, dof = "2", "pendulum_order25_events_trajectory_steps_per_s", 1685, 1 << 20, 25, 2

    def __init__(self, a):
        super().__init__(a)
        from taylorintegration_b200 import problems as P
        self.tape = P.pendulum_tape(events=True)
        self.q0 = P.pendulum_ensemble(self.N)
        self.t0, self.tmax, self.maxsteps, self.cap = 0.0, 100.0, 5000, 64
        self.kind, self.params, self.per = 1, None, 0

    def describe(self):
        return (f"pendulum! + event g = x[2] (test/rootfinding.jl:20-26), order {self.order}, abstol 1e-20, theta_i uniform(0.1, 3.1) "
                f"seed 20261104, tmax 100, N = {self.N} per GPU, event capacity {self.cap}, output = final state + events")

    def oracle_run(self, idx, nthreads):
        import oracle
        s = oracle.integrate(self.tape, self.order, np.ascontiguousarray(self.q0[:, idx]), self.t0, self.tmax, kind=1,
                             abstol=1e-20, maxsteps=self.maxsteps, event_capacity=self.cap, nthreads=nthreads)
        return _oracle_pack(s, ("nevents", "tevents", "xevents", "gresids"))


class Lorenz(Workload):
    key, metric, flops_per_step, default_n, order, dof = "3", "lorenz_lyapunov_order28_trajectory_steps_per_s", 24700, 1 << 16, 28, 3

    def __init__(self, a):
        super().__init__(a)
        from taylorintegration_b200 import problems as P
        self.tape = P.lorenz_tape()
        self.q0, self.params = P.lorenz_ensemble(self.N)
        self.t0, self.tmax, self.maxsteps = 0.0, 100.0, 20_000
        self.kind, self.per = 2, 1

    def describe(self):
        return (f"lorenz! Lyapunov spectrum (benchmark/Lyap_benchmark.jl), sigma 16, beta 4, rho_i = 28 + 64 i/(N-1), q0 = [19,20,50], "
                f"order {self.order}, abstol 1e-20, tmax 100, N = {self.N} per GPU, output = final state + exponents")

    def oracle_run(self, idx, nthreads):
        import oracle
        s = oracle.integrate(self.tape, self.order, np.ascontiguousarray(self.q0[:, idx]), self.t0, self.tmax,
                             np.ascontiguousarray(self.params[:, idx]), kind=2, abstol=1e-20, maxsteps=self.maxsteps,
                             nthreads=nthreads)
        return _oracle_pack(s, ("lam",))

    def min_sample(self, nthreads):
        return nthreads


class ManySpin(Workload):
    key, metric, flops_per_step, default_n, order, dof = "4", "manyspin100_order25_trajectory_steps_per_s", 6.5e5, 1 << 14, 25, 300

    def __init__(self, a):
        super().__init__(a)
        from taylorintegration_b200 import problems as P
        self.nn = 100
        self.tape = P.manyspin_tape(self.nn)
        self.q0 = P.manyspin_ensemble(self.nn, self.N)
        self.t0, self.tmax, self.maxsteps = 0.0, 10.0, 5000
        self.kind, self.params, self.per = 0, None, 0

    def describe(self):
        return (f"spin_odes! (benchmark/manyspin_benchmarks.jl:10-41), {self.nn} spins = 300 components, J = fix_J(100, 1.4), seeded B, "
                f"order {self.order}, abstol 1e-20, tmax 10, N = {self.N} members per GPU, output = final state")

    def oracle_run(self, idx, nthreads):
        import oracle
        s = oracle.integrate(self.tape, self.order, np.ascontiguousarray(self.q0[:, idx]), self.t0, self.tmax, abstol=1e-20,
                             maxsteps=self.maxsteps, nthreads=nthreads)
        return _oracle_pack(s)

    def min_sample(self, nthreads):
        return nthreads


class Jet(Workload):
    """Jet transport (TaylorN coefficients): 5a pendulum in 2 variables, 5b Kepler in 4 variables, polynomial order 8."""
    order = 28

    def __init__(self, a, which):
        self.key = which
        if which == "5a":
            self.metric, self.default_n, self.nvars, self.dof = "jet_transport_pendulum_n2_d8_trajectory_steps_per_s", 1 << 16, 2, 2
        else:
            # BASELINE: 2^16 neighbourhoods "sharded across 8 B200" = 8192 per GPU; the default one-GPU line uses 2048
            # (20 s per pass at ~1.9e4 steps/s) so that W + K passes stay within minutes; --ntraj 8192 is the per-GPU shard
            self.metric, self.default_n, self.nvars, self.dof = "jet_transport_kepler_n4_d8_trajectory_steps_per_s", 2048, 4, 4
        super().__init__(a)
        from taylorintegration_b200 import problems as P
        import taylorintegration_b200 as tb
        self.varorder = 8
        self.ncoef = math.comb(self.nvars + self.varorder, self.varorder)
        pairs = math.comb(2 * self.nvars + self.varorder, self.varorder)
        if which == "5a":
            self.tape = P.pendulum_tape(events=False)
            th = np.random.Generator(np.random.PCG64(20261104)).uniform(0.1, 3.0, self.N)
            th[0] = 1.3
            nominal = np.stack([th, np.zeros(self.N)])
            self.params, self.per = None, 0
            self.trange = np.arange(0, 49) * (T_PEND / 8)  # 0 : T/8 : 6T (docs/src/pendulum.md:92-106)
            self.t0, self.tmax, self.maxsteps = 0.0, float(self.trange[-1]), 2000
            self.flops_per_step = 2.0 * pairs * 2 * sum(range(1, 28))  # sincos!: 2 polynomial products per (k, i) pair
        else:
            self.tape = P.kepler_tape()
            nominal, self.params, _ = P.kepler_ensemble(self.N)
            self.per = 0
            self.trange = None
            self.t0, self.tmax, self.maxsteps = 0.0, 2 * np.pi, 5000
            prods = sum((k // 2 + 1) * 2 + k + 2 * k for k in range(28))  # 2 sqr!, pow!, 2 div! polynomial products per order
            self.flops_per_step = 2.0 * pairs * prods
        self.jt = tb.JetTransport(self.tape, self.order, self.nvars, self.varorder,
                                  tb._lib.MODE_FAST if a.mode == "fast" else tb._lib.MODE_PARITY)
        self.q0 = self.jt.initial(nominal)  # [dof, ncoef, N]
        self.kind = 0

    def describe(self):
        what = ("pendulum, q0 = [theta_i, 0] + xi, trange 0:T/8:6T" if self.key == "5a"
                else "kepler1!, q0(e_i) + xi, one period")
        return (f"jet transport {what}: TaylorN in {self.nvars} variables of order {self.varorder} ({self.ncoef} coefficients per "
                f"component), time order {self.order}, abstol 1e-20, N = {self.N} nominal points per GPU")

    def oracle_run(self, idx, nthreads):
        import oracle
        r = oracle.integrate_tn(self.tape, self.order, self.nvars, self.varorder, np.ascontiguousarray(self.q0[:, :, idx]),
                                self.t0, self.tmax, self.params, trange=self.trange, abstol=1e-20, maxsteps=self.maxsteps)
        return r

    def min_sample(self, nthreads):
        return 1 if self.key == "5b" else 4


def make_workload(a) -> Workload:
    return {"1": Kepler, "2": Pendulum, "3": Lorenz, "4": ManySpin}[a.config](a) if a.config in "1234" else Jet(a, a.config)


def config_dict(a, w, extra=None):
    c = {"workload": w.describe(), "config": a.config, "ntraj_per_gpu": w.N, "order": w.order, "mode": a.mode,
         "flops_per_trajectory_step": w.flops_per_step,
         "l2": "L2 flushed between timed iterations (256 MiB write)"}
    if a.config == "1":
        c["periods"] = a.periods
    if extra:
        c.update(extra)
    return c


# ---- CPU arm: the oracle (restatement of the reference) on all host threads ------------------------------------
def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = make_workload(a)
    per_step = max(2.0, min(20.0, 120.0 / max(1, a.steps + a.warmup)))
    rates, desc, cores = [], "", 1
    for i in range(a.warmup + a.steps):
        r, cores, desc, _, _ = w.cpu_sample(per_step)
        if i >= a.warmup:
            rates.append(r)
    v = float(np.mean(rates))
    line = {"impl": "reference", "metric": w.metric, "value": v,
            "unit": "trajectory-steps/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": None, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(a, w),
            "cpu_baseline": {"value": v, "unit": "trajectory-steps/s", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": "trajectory-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference is pure Julia (no julia binary here): CPU arm = oracle/tb_oracle.cpp (tb_tn_oracle.cpp for "
                    "jet transport), the C++ restatement of taylorinteg, -O2 -ffp-contract=off, one OS thread per host core"}
    print(json.dumps(line), flush=True)


# ---- GPU arm ------------------------------------------------------------------------------------------------------
def parity_check(w, idx, ref, got) -> dict:
    """Bit-for-bit comparison of the oracle sample with the GPU results at the same trajectory indices."""
    fields, ok = [], True
    for f, r in ref.items():
        if f not in got or got[f] is None:
            continue
        g = got[f][..., idx]
        if f in ("tevents", "gresids", "xevents"):  # valid entries only
            ne = ref["nevents"]
            m = np.arange(r.shape[-2])[:, None] < ne[None, :]
            same = np.array_equal(g[..., m], r[..., m])
        else:
            same = np.array_equal(g, r, equal_nan=True)
        fields.append(f)
        ok = ok and bool(same)
    return {"n": int(idx.size), "bit_exact": bool(ok), "fields": fields,
            "what": "GPU results vs oracle on the cpu_baseline sample of the same ensemble (array_equal)"}


def run_b200(a):
    import torch
    import torch.distributed as dist
    import taylorintegration_b200 as tb
    from taylorintegration_b200 import _lib as L
    from taylorintegration_b200.api import _column_tape
    from taylorintegration_b200.sharding import auto_block, shard_indices
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    lib = L.lib()
    L.check(lib.tb_set_device(local), "tb_set_device")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    w = make_workload(a)          # the whole ensemble of this rank (weak) or of the job (strong)
    strong = a.scaling == "strong" and world > 1
    Nglob = w.N
    own = shard_indices(Nglob, world, rank, auto_block(Nglob, world)) if strong else np.arange(Nglob)
    N = int(own.size)
    mode = L.MODE_PARITY if a.mode == "parity" else L.MODE_FAST
    dof = w.dof
    jet = isinstance(w, Jet)
    wide = isinstance(w, ManySpin)
    q_loc = np.ascontiguousarray(w.q0[..., own])
    p_loc = w.params if (w.params is None or not w.per) else np.ascontiguousarray(w.params[:, own])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    # ---- result buffers (pinned host for the e2e arm, device for the resident arm) ----
    def alloc(shape, dtype, pinned):
        return torch.empty(shape, dtype=dtype).pin_memory() if pinned else torch.empty(shape, dtype=dtype, device=dev)

    def outputs(pinned):
        o = {"xfinal": alloc((dof, w.ncoef, N) if jet else (dof, N), torch.float64, pinned),
             "tfinal": alloc((N,), torch.float64, pinned), "nsteps": alloc((N,), torch.int64, pinned),
             "status": alloc((N,), torch.int32, pinned)}
        if w.kind == 1:
            o.update(tevents=alloc((w.cap, N), torch.float64, pinned), xevents=alloc((dof, w.cap, N), torch.float64, pinned),
                     gresids=alloc((w.cap, N), torch.float64, pinned), nevents=alloc((N,), torch.int32, pinned))
        if w.kind == 2:
            o["lam"] = alloc((dof, N), torch.float64, pinned)
        if jet and w.trange is not None:
            o["x_trange"] = alloc((dof, w.ncoef, w.trange.size, N), torch.float64, pinned)
        return o

    def out_struct(o):
        s = L.TbOut()
        s.xfinal, s.tfinal = C.cast(o["xfinal"].data_ptr(), L._DP), C.cast(o["tfinal"].data_ptr(), L._DP)
        s.nsteps = C.cast(o["nsteps"].data_ptr(), C.POINTER(C.c_int64))
        s.status = C.cast(o["status"].data_ptr(), C.POINTER(C.c_int32))
        if "tevents" in o:
            s.tevents, s.gresids = C.cast(o["tevents"].data_ptr(), L._DP), C.cast(o["gresids"].data_ptr(), L._DP)
            s.xevents = C.cast(o["xevents"].data_ptr(), L._DP)
            s.nevents = C.cast(o["nevents"].data_ptr(), C.POINTER(C.c_int32))
        if "lam" in o:
            s.lambda_ = C.cast(o["lam"].data_ptr(), L._DP)
        return s

    h_q0 = torch.from_numpy(q_loc).pin_memory()
    h_par = None if p_loc is None else torch.from_numpy(np.ascontiguousarray(p_loc, dtype=np.float64)).pin_memory()
    h_o = outputs(True)
    # config 1: the ensemble is cost-ordered (steps per period grow with the eccentricity), so the scheduling hint
    # `sort_by_cost` lets the work queue hand out the expensive end first (same results, shorter tail)
    queue_reverse = a.queue == "reverse" or (a.queue == "auto" and a.config in ("1", "2", "3"))
    opts = tb.make_opts(abstol=1e-20, maxsteps=w.maxsteps, params_per_traj=w.per,
                        event_capacity=getattr(w, "cap", 0), sort_by_cost=int(queue_reverse))
    kernel_ms = []          # device time of the fused kernel of every host-buffer call (tb_last_kernel_ms)
    info, khash, kname = {}, None, "tb_ensemble_kernel"
    gathered = None
    if jet or wide:
        # the TaylorN and large-state programs have host-buffer entry points only: `value` is the device time of their
        # fused kernel (CUDA events inside the library, inputs resident at that point), `e2e` the whole call
        d_o = None
        if wide:
            from taylorintegration_b200.api import get_wide_kernel
            k = get_wide_kernel(w.tape, w.order, mode)
            kname = "tb_wide_kernel"
            info = k.info()
            h_out = out_struct(h_o)

            def step_host():
                L.check(lib.tb_wide_integrate(k.handle, N, h_q0.data_ptr(), None, w.t0, w.tmax, None, 0, C.byref(opts),
                                              C.byref(h_out)), "tb_wide_integrate")
                kernel_ms.append(lib.tb_last_kernel_ms())
        else:
            kname = "tb_tn_kernel"
            tn = L.TbTnOut()
            tn.xfinal, tn.tfinal = C.cast(h_o["xfinal"].data_ptr(), L._DP), C.cast(h_o["tfinal"].data_ptr(), L._DP)
            tn.nsteps = C.cast(h_o["nsteps"].data_ptr(), C.POINTER(C.c_int64))
            tn.status = C.cast(h_o["status"].data_ptr(), C.POINTER(C.c_int32))
            h_tr = None
            if w.trange is not None:
                tn.xv_trange = C.cast(h_o["x_trange"].data_ptr(), L._DP)
                h_tr = np.ascontiguousarray(w.trange)

            def step_host():
                L.check(lib.tb_tn_integrate(w.jt._t, N, h_q0.data_ptr(), None if h_par is None else h_par.data_ptr(), w.t0,
                                            w.tmax, None if h_tr is None else h_tr.ctypes.data, 0 if h_tr is None else h_tr.size,
                                            C.byref(opts), C.byref(tn)), "tb_tn_integrate")
                kernel_ms.append(lib.tb_last_kernel_ms())
        step_device = None
    else:
        ktape, kkind, kw = w.tape, w.kind, dict(flags=a.kflags, reg_budget=a.reg_budget, block_threads=a.block_threads)
        if w.kind == 0 and not a.full_outputs_kernel:
            kw["flags"] |= 262144  # TB_CF_FINAL_ONLY: the build the API picks for output="final"
        if w.kind == 2:
            ktape, kkind = _column_tape(w.tape), L.KIND_LYAPCOL
        k = tb.Kernel(ktape, w.order, kkind, mode, **kw)
        info = k.info()
        khash = hashlib.sha1(k.source.encode()).hexdigest()[:16]
        d_q0 = h_q0.to(dev)
        d_par = None if h_par is None else h_par.to(dev)
        d_o = outputs(False)
        d_out, h_out = out_struct(d_o), out_struct(h_o)
        d_work = torch.zeros(int(lib.tb_work_bytes()), dtype=torch.uint8, device=dev)
        if world > 1:
            gathered = torch.empty((world,) + tuple(d_o["xfinal"].shape), dtype=torch.float64, device=dev)

        def step_device(kk=None):
            L.check(lib.tb_integrate_device((kk or k).handle, N, d_q0.data_ptr(), None if d_par is None else d_par.data_ptr(),
                                            w.t0, w.tmax, None, 0, C.byref(opts), C.byref(d_out), d_work.data_ptr(),
                                            stream.cuda_stream), "tb_integrate_device")
            if world > 1:  # the path's only exchange: final gather of the shards' results
                dist.all_gather_into_tensor(gathered, d_o["xfinal"])

        def step_host():
            L.check(lib.tb_integrate(k.handle, N, h_q0.data_ptr(), None if h_par is None else h_par.data_ptr(), w.t0, w.tmax,
                                     C.byref(opts), C.byref(h_out)), "tb_integrate")
            kernel_ms.append(lib.tb_last_kernel_ms())
            if world > 1:  # the e2e arm carries the gather as well (from the host results)
                d_o["xfinal"].copy_(h_o["xfinal"], non_blocking=True)
                dist.all_gather_into_tensor(gathered, d_o["xfinal"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, events: bool):
        """W warm-ups, then K timed steps; returns (seconds, max over ranks; per-step device ms)."""
        for _ in range(a.warmup):
            fn()
        barrier()
        kms, total = [], 0.0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(a.steps):
            flush.fill_(1)  # L2 flush between timed iterations
            barrier()
            t0 = time.perf_counter()
            if events:
                e0.record(stream)
            fn()
            if events:
                e1.record(stream)
            barrier()
            total += time.perf_counter() - t0
            if events:
                kms.append(e0.elapsed_time(e1))
        if events:  # device time is the clock for the resident arm
            total = sum(kms) / 1e3
        t = torch.tensor([total], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), kms

    peak_tf, _ = tb.fp64_peak(0.3)
    clocks = ClockSampler(local)
    clocks.start()
    launches0 = lib.tb_launch_count()
    if step_device is not None:
        secs, kms = timed(step_device, True)
        res = d_o
    else:
        kernel_ms.clear()
        secs_wall, _ = timed(step_host, False)
        kms = kernel_ms[-a.steps:]
        t = torch.tensor([sum(kms) / 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
        res = h_o
    launches = (lib.tb_launch_count() - launches0) // (a.warmup + a.steps) * a.steps
    clk = clocks.stop()
    steps_rank = torch.tensor([float(res["nsteps"].sum().item())], dtype=torch.float64, device=dev)
    ok = bool((res["status"] == 0).all().item()) and bool((res["tfinal"] == w.tmax).all().item())
    if world > 1:
        dist.all_reduce(steps_rank, op=dist.ReduceOp.SUM)
    total_steps = float(steps_rank.item())
    value = total_steps * a.steps / secs
    got = {f: v.cpu().numpy() for f, v in res.items()}
    e2e = None
    if not a.no_e2e:
        if step_device is not None:
            secs_h, _ = timed(step_host, False)
            assert torch.equal(h_o["nsteps"], d_o["nsteps"].cpu()) and torch.equal(h_o["xfinal"], d_o["xfinal"].cpu()), \
                "host-buffer path != device path"
        else:
            secs_h = secs_wall
        h2d = int(h_q0.numel() * 8 + (0 if h_par is None else h_par.numel() * 8))
        d2h = int(sum(v.numel() * v.element_size() for v in h_o.values()))
        if w.kind == 1:  # event buffers: only the rows in use travel (max nevents of capacity)
            mx = int(h_o["nevents"].max().item())
            d2h -= int((w.cap - mx) * N * 8 * (2 + dof))
        e2e = {"value": total_steps * a.steps / secs_h, "unit": "trajectory-steps/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h,
               "api": ("tb_wide_integrate" if wide else "tb_tn_integrate" if jet else "tb_integrate") + " (C ABI, pinned host buffers)"}
    fast = None
    if a.config == "1" and a.mode == "parity" and not a.no_fast:
        # the same ensemble through the FMA build (TB_MODE_FAST): reported beside the bit-exact headline
        x_par = d_o["xfinal"].clone()
        kf = tb.Kernel(w.tape, w.order, L.KIND_PLAIN, L.MODE_FAST, flags=0 if a.full_outputs_kernel else 262144)
        secs_f, kms_f = timed(lambda: step_device(kf), True)
        tsf = torch.tensor([float(d_o["nsteps"].sum().item())], dtype=torch.float64, device=dev)
        rel = ((d_o["xfinal"] - x_par).abs().amax(0) / x_par.abs().amax(0)).max().reshape(1)
        if world > 1:
            dist.all_reduce(tsf, op=dist.ReduceOp.SUM)
            dist.all_reduce(rel, op=dist.ReduceOp.MAX)
        fast = {"value": float(tsf.item()) * a.steps / secs_f, "unit": "trajectory-steps/s",
                "ms_per_step": secs_f / a.steps * 1e3,
                "roofline_frac": float(tsf.item()) / world * w.flops_per_step / (float(np.mean(kms_f)) * 1e-3) / 1e12 / peak_tf,
                "max_rel_diff_final_state_vs_parity": float(rel.item()),
                "note": "TB_MODE_FAST: FMA + re-associated sums; not bit-identical to the reference order"}
    rc = 0
    if rank == 0:
        ms_kernel = float(np.mean(kms))
        # DRAM traffic of one launch of THIS kernel on this workload, from the committed `ncu --set full` capture
        # (profiles/ncu_traffic.json, stamped with the sha1 of the generated kernel source); null when it does not match
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                t = json.load(f)
            if khash is not None and t.get("kernel_source_sha1") == khash and \
                    (t.get("config"), t.get("ntraj"), t.get("mode")) == (a.config, N, a.mode):
                traffic = t["dram_bytes_per_launch"]
        except (OSError, ValueError, KeyError):
            pass
        ach_tf = (total_steps / world) * w.flops_per_step / (ms_kernel * 1e-3) / 1e12
        alg_bytes = int(q_loc.nbytes + (0 if p_loc is None else np.asarray(p_loc).nbytes) +
                        sum(v.numel() * v.element_size() for f, v in h_o.items() if f not in ("tevents", "xevents", "gresids")))
        line = {"metric": w.metric, "value": value, "unit": "trajectory-steps/s",
                "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": secs / a.steps * 1e3,
                "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": config_dict(a, w, {"queue": "reverse (tb_opts_t.sort_by_cost = 1: the ensemble's cost grows with the index, the most expensive "
                                                      "trajectories are handed out first; results do not depend on it)" if queue_reverse else "forward",
                                             "trajectory_steps_per_pass": total_steps, "all_ok": ok, "kernel": info,
                                              "kernel_source_sha1": khash, "ntraj_total": Nglob * (1 if strong else world)}),
                "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
                "roofline": {"bound": "fp64", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s",
                             "frac": ach_tf / peak_tf, "traffic": traffic,
                             "algorithmic_bytes_per_launch": alg_bytes,
                             "peak_source": "tb_fp64_peak: register-only DFMA chains measured in this run "
                                            "(MEASURED_PEAKS.json has no FP64 entry; nominal 37.2 TFLOP/s)",
                             "kernel": kname, "kernel_ms": ms_kernel,
                             "pct_of_fp64_peak": 100.0 * ach_tf / peak_tf}}
        if fast:
            line["fast_mode"] = fast
        if not a.no_cpu_baseline:
            r, cores, desc, idx, ref = w.cpu_sample(a.cpu_seconds)
            line["cpu_baseline"] = {"value": r, "unit": "trajectory-steps/s", "cores": cores, "kind": "port", "sample": desc}
            if a.mode == "parity":
                if strong:  # rank 0 holds the round-robin blocks `own`: compare the sample members it owns
                    pos = {int(g): i for i, g in enumerate(own)}
                    keep = np.array([j for j, g in enumerate(idx) if int(g) in pos], dtype=np.int64)
                    ref = {f: v[..., keep] for f, v in ref.items()}
                    idx = np.array([pos[int(idx[j])] for j in keep], dtype=np.int64)
                line["parity_sample"] = parity_check(w, idx, ref, got)
                if not line["parity_sample"]["bit_exact"] or not ok:
                    rc = 3
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if rc:
        raise SystemExit(rc)


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
