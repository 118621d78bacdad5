"""Large-state path (csrc/tb_wide.cu, one CTA per trajectory): the many-spin system of
benchmark/manyspin_benchmarks.jl and op coverage through the same interpreter."""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import api, problems as P


def test_fusion_pass_statistics():
    """The inner `for a = 1:nn` loop (manyspin_benchmarks.jl:27-31) collapses to one fused entry per spin."""
    for nn in (3, 7, 100):
        tape = P.manyspin_tape(nn)
        assert len(tape.ops) == 2 * nn * nn + 6 * nn  # nn^2 mul + nn^2 add + per spin: neg, 3 mul, add, (dq ops)
        info = api.WideKernel(tape, 25).info()
        assert info["entries"] == 6 * nn and info["levels"] == 3 and info["max_width"] == 3 * nn
    assert api.is_wide(P.manyspin_tape(7)) and not api.is_wide(P.kepler_tape())


def test_oracle_manyspin_invariants():
    """|s_i| = 1 for every spin and the energy are conserved (reference config: nn = 7, order 26)."""
    nn = 7
    B, J = P.fix_B(nn), P.fix_J(nn)
    tape = P.manyspin_tape(nn, B, J)
    q0 = P.manyspin_ensemble(nn, 3)
    s = oracle.integrate(tape, 26, q0, 0.0, 20.0, maxsteps=5000)
    assert (s.status == 0).all() and (s.tfinal == 20.0).all()
    v = s.xfinal.reshape(nn, 3, -1)
    assert np.abs((v ** 2).sum(1) - 1.0).max() < 1e-13

    def energy(q):  # H = sum_i B_i z_i - 1/2 x^T J x is conserved by spin_odes!
        v = q.reshape(nn, 3, -1)
        return (B[:, None] * v[:, 2]).sum(0) - 0.5 * np.einsum("ai,an,in->n", J, v[:, 0], v[:, 0])

    assert np.abs(energy(s.xfinal) - energy(q0)).max() < 1e-12


@pytest.mark.gpu
def test_gpu_manyspin7_bitwise():
    tape = P.manyspin_tape(7)
    q0 = P.manyspin_ensemble(7, 40)
    sol = tb.taylorinteg(tape, q0, 0.0, 10.0, 26, 1e-20, maxsteps=5000)
    ref = oracle.integrate(tape, 26, q0, 0.0, 10.0, maxsteps=5000)
    assert np.array_equal(sol.nsteps, ref.nsteps) and np.array_equal(sol.tfinal, ref.tfinal) and (sol.status == 0).all()
    assert np.array_equal(sol.xfinal, ref.xfinal)


@pytest.mark.gpu
def test_gpu_manyspin100_bitwise_and_trange():
    """Config 4 shape: 100 spins (300 components), order 25, abstol 1e-20; short horizon for the oracle."""
    tape = P.manyspin_tape(100)
    q0 = P.manyspin_ensemble(100, 3)
    sol = tb.taylorinteg(tape, q0, 0.0, 0.5, 25, 1e-20, maxsteps=5000)
    ref = oracle.integrate(tape, 25, q0, 0.0, 0.5, maxsteps=5000)
    assert np.array_equal(sol.nsteps, ref.nsteps) and (sol.status == 0).all()
    assert np.array_equal(sol.xfinal, ref.xfinal)
    assert np.abs((sol.xfinal.reshape(100, 3, -1) ** 2).sum(1) - 1.0).max() < 1e-13
    tr = np.linspace(0.0, 0.3, 7)
    st = tb.taylorinteg(tape, q0, tr, 25, 1e-20, maxsteps=5000)
    rt = oracle.integrate(tape, 25, q0, trange=tr, maxsteps=5000)
    assert np.array_equal(st.x_trange, rt.x_trange, equal_nan=True)


@pytest.mark.gpu
def test_gpu_wide_interpreter_matches_thread_kernel():
    """Kepler (sqr, pow, div) and pendulum (sincos) forced through the wide interpreter == oracle == thread kernel."""
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(33)
    tmax = 2 * np.pi * 2
    a = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=2000, kernel_opts={"wide": True})
    b = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=2000)
    ref = oracle.integrate(tape, 28, q0, 0.0, tmax, mu, maxsteps=2000)
    assert np.array_equal(a.nsteps, ref.nsteps) and np.array_equal(a.xfinal, ref.xfinal)
    assert np.array_equal(a.xfinal, b.xfinal)
    tp = P.pendulum_tape(events=False)
    qp = P.pendulum_ensemble(17)
    qp = qp[0] if isinstance(qp, tuple) else qp
    c = tb.taylorinteg(tp, qp, 0.0, 20.0, 25, 1e-20, maxsteps=2000, kernel_opts={"wide": True})
    rp = oracle.integrate(tp, 25, qp, 0.0, 20.0, maxsteps=2000)
    assert np.array_equal(c.nsteps, rp.nsteps) and np.array_equal(c.xfinal, rp.xfinal)
    f = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=2000, kernel_opts={"wide": True}, mode="fast")
    assert np.abs(f.xfinal - ref.xfinal).max() < 1e-9


def test_cluster_layout_of_the_shared_memory_kernel():
    """tb_wide2.cu: many-spin 100 = one coupling cluster per spin (+ its small dq[3i-2] cluster on the same thread),
    400 stored series (300 state jets + 100 coupling sums) x 26 doubles = 83 KB in shared memory, two CTAs per SM (the
    100 x 100 coupling constants stay in L1/L2: measured faster than 1 CTA per SM with them in shared memory); the two kinds of cluster (coupling, dq[3i-2]) each share one copy of their entries (group programs)."""
    info = api.WideKernel(P.manyspin_tape(100), 25).info()
    assert info["threads"] == 128 and info["stored_series"] == 400 and info["scalar_slots"] <= 4
    assert info["smem_bytes"] <= 110 * 1024 and info["ctas_per_sm"] == 2 and info["groups"] == 2
    # the two group programs are specialised by NVRTC (straight-line entries, the head of the coupling constants in
    # registers, the rest through L1/L2): no spills
    wk = api.WideKernel(P.manyspin_tape(100), 25)
    assert info["generated"] == 1 and info["local_bytes"] == 0 and info["coupling_in_smem"] == 0
    assert "tb_w2_groups" in wk.source and "tb_w2_kernel" in wk.log
    # the two Cauchy products of a spin (z_i * S_i, -y_i * S_i) share S_i: fused into one W_MUL2 entry (op 102), the
    # second product moved up past the one-term coupling entry that does not touch it
    assert wk.source.count("o.op = 102;") == 1 and wk.source.count("o.op = 5;") == 1  # 5: the scalar * series of group 1
    k = api.WideKernel(P.kepler_tape(), 28).info()   # one cluster: everything depends on r^3
    assert k["threads"] == 32 and k["stored_series"] >= 8


@pytest.mark.gpu
def test_gpu_manyspin100_config4_sample_and_kernel_equivalence(monkeypatch):
    """BASELINE config 4 (100 spins, order 25, tmax 10): 64 members on the shared-memory cluster kernel, a strided
    sample of them bit for bit against the oracle, all of them against the global-workspace kernel (TB_WIDE_V1)."""
    tape = P.manyspin_tape(100)
    q0 = P.manyspin_ensemble(100, 64)
    sol = tb.taylorinteg(tape, q0, 0.0, 10.0, 25, 1e-20, maxsteps=5000)
    assert (sol.status == 0).all() and (sol.tfinal == 10.0).all()
    assert np.abs((sol.xfinal.reshape(100, 3, -1) ** 2).sum(1) - 1.0).max() < 1e-12
    idx = np.arange(0, 64, 8)
    ref = oracle.integrate(tape, 25, np.ascontiguousarray(q0[:, idx]), 0.0, 10.0, maxsteps=5000, nthreads=8)
    assert np.array_equal(sol.nsteps[idx], ref.nsteps) and np.array_equal(sol.xfinal[:, idx], ref.xfinal)
    monkeypatch.setenv("TB_WIDE_V1", "1")
    v1 = tb.taylorinteg(tape, q0, 0.0, 10.0, 25, 1e-20, maxsteps=5000)
    assert np.array_equal(v1.nsteps, sol.nsteps) and np.array_equal(v1.xfinal, sol.xfinal)
    f = tb.taylorinteg(tape, q0[:, :8], 0.0, 2.0, 25, 1e-20, maxsteps=5000, mode="fast")
    monkeypatch.delenv("TB_WIDE_V1")
    f2 = tb.taylorinteg(tape, q0[:, :8], 0.0, 2.0, 25, 1e-20, maxsteps=5000, mode="fast")
    assert np.abs(f.xfinal - f2.xfinal).max() < 1e-9
