#!/usr/bin/env python3
"""bench.py -- Kepler order-28 ensemble throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            # this framework (libtaylorb200.so, sm_100a)
    python bench.py --impl reference --gpus N --steps K ...  # the CPU restatement of the reference on host cores

One "step" = one pass of the hot path over one batch: the whole ensemble (default 2^20 eccentricities per GPU,
order 28, abstol 1e-20, 100 periods, final-state output) integrated from t0 to tmax in ONE kernel launch.
`value` = trajectory-steps/s with q0 already resident in HBM; `e2e` = the same through the host-buffer C-ABI
call tb_integrate (pinned host q0 -> device, kernel, results -> host inside the timed region).
Multi-GPU: one process per GPU (torchrun), ensembles shard with no data-path collective (weak scaling); the
only communication is the final NCCL all-gather of the per-rank results, inside the timed region.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

KEPLER_FLOPS_PER_STEP = 4733  # jet 4493 + Horner 224 + stepsize 16 (order 28)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ntraj", type=int, default=1 << 20, help="trajectories per GPU")
    ap.add_argument("--periods", type=float, default=100.0)
    ap.add_argument("--order", type=int, default=28)
    ap.add_argument("--mode", default="parity", choices=["parity", "fast"])
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fast", action="store_true", help="skip the extra fast-mode (FMA) measurement")
    ap.add_argument("--reg-budget", type=int, default=0, help="doubles of jet storage in registers (kernel variant 2)")
    ap.add_argument("--block-threads", type=int, default=0)
    ap.add_argument("--kflags", type=int, default=0, help="TB_CF_* compile flags (4 = rolled variant 1)")
    ap.add_argument("--full-outputs-kernel", action="store_true",
                    help="use the build that also carries trange / per-step / dense output code (default: the "
                         "final-state build, TB_CF_FINAL_ONLY, which is what the API picks for output='final')")
    return ap.parse_args()


def workload(a):
    from taylorintegration_b200 import problems as P
    q0, mu, e = P.kepler_ensemble(a.ntraj)
    return P.kepler_tape(), q0, mu, dict(t0=0.0, tmax=2 * np.pi * a.periods, abstol=1e-20, maxsteps=10_000)


def config_dict(a, extra=None):
    c = {"workload": f"kepler1! planar 2-body, order {a.order}, abstol 1e-20, e_i = 0.9*i/(N-1), "
                     f"{a.periods:g} periods, N = {a.ntraj} trajectories per GPU, output = final state + nsteps",
         "ntraj_per_gpu": a.ntraj, "order": a.order, "periods": a.periods, "mode": a.mode,
         "flops_per_trajectory_step": KEPLER_FLOPS_PER_STEP,
         "l2": "L2 flushed between timed iterations (256 MiB write); inputs are 32 MiB per GPU"}
    if extra:
        c.update(extra)
    return c


# ---- CPU arm: the oracle (restatement of the reference) on all host threads ------------------------------------
def cpu_sample(a, target_seconds, nthreads=None):
    """Times the oracle on a strided sample of the same ensemble; returns (steps/s, cores, description, seconds)."""
    import oracle
    tape, q0, mu, w = workload(a)
    nthreads = nthreads or os.cpu_count() or 1
    n = min(a.ntraj, 4 * nthreads)
    rate, secs, used = 0.0, 0.0, 0
    for _ in range(3):
        idx = np.linspace(0, a.ntraj - 1, n).round().astype(np.int64)
        q = np.ascontiguousarray(q0[:, idx])
        t0 = time.perf_counter()
        s = oracle.integrate(tape, a.order, q, w["t0"], w["tmax"], mu, abstol=w["abstol"], maxsteps=w["maxsteps"],
                             nthreads=nthreads)
        secs = time.perf_counter() - t0
        steps = int(s.nsteps.sum())
        rate, used = steps / secs, n
        if secs >= 0.5 * target_seconds or n >= a.ntraj:
            break
        n = int(min(a.ntraj, max(n * 2, n * target_seconds / max(secs, 1e-3) * 0.9)))
    desc = (f"{used} of {a.ntraj} trajectories (evenly strided over e), full {a.periods:g} periods each, "
            f"{nthreads} threads, {secs:.1f} s")
    return rate, nthreads, desc, secs


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step = max(2.0, min(20.0, 120.0 / max(1, a.steps + a.warmup)))
    rates, desc, cores = [], "", 1
    for i in range(a.warmup + a.steps):
        r, cores, desc, _ = cpu_sample(a, per_step)
        if i >= a.warmup:
            rates.append(r)
    v = float(np.mean(rates))
    line = {"impl": "reference", "metric": "kepler_order28_ensemble_trajectory_steps_per_s", "value": v,
            "unit": "trajectory-steps/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(a),
            "cpu_baseline": {"value": v, "unit": "trajectory-steps/s", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": "trajectory-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference is pure Julia (no julia binary here): CPU arm = oracle/tb_oracle.cpp, the C++ "
                    "restatement of taylorinteg, -O2 -ffp-contract=off, one OS thread per host core"}
    print(json.dumps(line), flush=True)


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


# ---- GPU arm ------------------------------------------------------------------------------------------------------
def run_b200(a):
    import torch
    import torch.distributed as dist
    import taylorintegration_b200 as tb
    from taylorintegration_b200 import _lib as L
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
    tape, q0, mu, w = workload(a)
    N, dof = a.ntraj, 4
    k = tb.Kernel(tape, a.order, L.KIND_PLAIN, L.MODE_PARITY if a.mode == "parity" else L.MODE_FAST,
                  reg_budget=a.reg_budget, block_threads=a.block_threads,
                  flags=a.kflags | (0 if a.full_outputs_kernel else 262144))
    opts = tb.make_opts(abstol=w["abstol"], maxsteps=w["maxsteps"])
    # pinned host buffers (e2e arm) and device-resident copies (kernel arm)
    h_q0 = torch.from_numpy(q0).pin_memory()
    h_mu = torch.from_numpy(mu).pin_memory()
    h_xf = torch.empty((dof, N), dtype=torch.float64).pin_memory()
    h_tf = torch.empty(N, dtype=torch.float64).pin_memory()
    h_ns = torch.empty(N, dtype=torch.int64).pin_memory()
    h_st = torch.empty(N, dtype=torch.int32).pin_memory()
    d_q0, d_mu = h_q0.to(dev), h_mu.to(dev)
    d_xf = torch.empty((dof, N), dtype=torch.float64, device=dev)
    d_tf = torch.empty(N, dtype=torch.float64, device=dev)
    d_ns = torch.empty(N, dtype=torch.int64, device=dev)
    d_st = torch.empty(N, dtype=torch.int32, device=dev)
    d_work = torch.zeros(int(lib.tb_work_bytes()), dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    gathered = torch.empty((world, dof, N), dtype=torch.float64, device=dev) if world > 1 else None

    def out_struct(xf, tf, ns, st):
        o = L.TbOut()
        o.xfinal = C.cast(xf.data_ptr(), L._DP)
        o.tfinal = C.cast(tf.data_ptr(), L._DP)
        o.nsteps = C.cast(ns.data_ptr(), C.POINTER(C.c_int64))
        o.status = C.cast(st.data_ptr(), C.POINTER(C.c_int32))
        return o

    d_out, h_out = out_struct(d_xf, d_tf, d_ns, d_st), out_struct(h_xf, h_tf, h_ns, h_st)
    stream = torch.cuda.current_stream()

    def step_device():
        L.check(lib.tb_integrate_device(k.handle, N, d_q0.data_ptr(), d_mu.data_ptr(), w["t0"], w["tmax"], None, 0,
                                        C.byref(opts), C.byref(d_out), d_work.data_ptr(), stream.cuda_stream),
                "tb_integrate_device")
        if world > 1:  # the path's only exchange: final gather of the shards' results
            dist.all_gather_into_tensor(gathered, d_xf)

    kcur = [k]

    def step_device_k():
        L.check(lib.tb_integrate_device(kcur[0].handle, N, d_q0.data_ptr(), d_mu.data_ptr(), w["t0"], w["tmax"], None, 0,
                                        C.byref(opts), C.byref(d_out), d_work.data_ptr(), stream.cuda_stream),
                "tb_integrate_device")

    def step_host():
        L.check(lib.tb_integrate(k.handle, N, h_q0.data_ptr(), h_mu.data_ptr(), w["t0"], w["tmax"], C.byref(opts),
                                 C.byref(h_out)), "tb_integrate")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, events: bool):
        """W warm-ups, then K timed steps; returns (seconds max over ranks, kernel ms list)."""
        for _ in range(a.warmup):
            fn()
        barrier()
        kms = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        total = 0.0
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
    secs, kms = timed(step_device, True)
    launches = (lib.tb_launch_count() - launches0) // (a.warmup + a.steps) * a.steps
    clk = clocks.stop()
    steps_rank = torch.tensor([float(d_ns.sum().item())], dtype=torch.float64, device=dev)
    ok = bool((d_st == 0).all().item()) and bool((d_tf == w["tmax"]).all().item())
    if world > 1:
        dist.all_reduce(steps_rank, op=dist.ReduceOp.SUM)
    total_steps = float(steps_rank.item())
    value = total_steps * a.steps / secs
    e2e = None
    if not a.no_e2e:
        secs_h, _ = timed(step_host, False)
        assert torch.equal(h_ns, d_ns.cpu()) and torch.equal(h_xf, d_xf.cpu()), "host-buffer path != device path"
        e2e = {"value": total_steps * a.steps / secs_h, "unit": "trajectory-steps/s",
               "h2d_bytes_per_step": int(h_q0.numel() * 8 + h_mu.numel() * 8),
               "d2h_bytes_per_step": int(h_xf.numel() * 8 + h_tf.numel() * 8 + h_ns.numel() * 8 + h_st.numel() * 4),
               "api": "tb_integrate (C ABI, pinned host buffers)"}
    fast = None
    if a.mode == "parity" and not a.no_fast:
        # the same ensemble through the FMA build (TB_MODE_FAST): reported beside the bit-exact headline
        x_par = d_xf.clone()
        kcur[0] = tb.Kernel(tape, a.order, L.KIND_PLAIN, L.MODE_FAST, flags=0 if a.full_outputs_kernel else 262144)
        secs_f, kms_f = timed(step_device_k, True)
        tsf = torch.tensor([float(d_ns.sum().item())], dtype=torch.float64, device=dev)
        rel = ((d_xf - x_par).abs().amax(0) / x_par.abs().amax(0)).max().reshape(1)
        if world > 1:
            dist.all_reduce(tsf, op=dist.ReduceOp.SUM)
            dist.all_reduce(rel, op=dist.ReduceOp.MAX)
        fast = {"value": float(tsf.item()) * a.steps / secs_f, "unit": "trajectory-steps/s",
                "ms_per_step": secs_f / a.steps * 1e3,
                "roofline_frac": float(tsf.item()) / world * KEPLER_FLOPS_PER_STEP / (float(np.mean(kms_f)) * 1e-3) / 1e12 / peak_tf,
                "max_rel_diff_final_state_vs_parity": float(rel.item()),
                "note": "TB_MODE_FAST: FMA + re-associated sums; not bit-identical to the reference order"}
    if rank != 0:
        return
    ms_kernel = float(np.mean(kms))
    # DRAM traffic of one launch of this kernel on this workload, from the committed `ncu --set full` capture
    # (profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum); null when no capture matches
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)
        if (t.get("ntraj"), t.get("periods"), t.get("order"), t.get("mode")) == (a.ntraj, a.periods, a.order, a.mode):
            traffic = t["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    ach_tf = (total_steps / world) * KEPLER_FLOPS_PER_STEP / (ms_kernel * 1e-3) / 1e12
    info = k.info()
    line = {"metric": "kepler_order28_ensemble_trajectory_steps_per_s", "value": value, "unit": "trajectory-steps/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": secs / a.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(a, {"trajectory_steps_per_pass": total_steps, "all_ok": ok, "kernel": info}),
            "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": {"bound": "fp64", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": ach_tf / peak_tf, "traffic": traffic,
                         "algorithmic_bytes_per_launch": int(N * (4 * 8 + 4 * 8 + 8 + 8 + 4) + 8),
                         "peak_source": "tb_fp64_peak: register-only DFMA chains measured in this run "
                                        "(MEASURED_PEAKS.json has no FP64 entry; nominal 37.2 TFLOP/s)",
                         "kernel": "tb_ensemble_kernel", "kernel_ms": ms_kernel,
                         "pct_of_fp64_peak": 100.0 * ach_tf / peak_tf}}
    if fast:
        line["fast_mode"] = fast
    if not a.no_cpu_baseline:
        r, cores, desc, _ = cpu_sample(a, a.cpu_seconds)
        line["cpu_baseline"] = {"value": r, "unit": "trajectory-steps/s", "cores": cores, "kind": "port", "sample": desc}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
