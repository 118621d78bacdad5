"""Committed golden fixtures (tests/golden/, made by tests/golden/make_golden.py).

CPU: the oracle reproduces the fixtures bit-for-bit (guards oracle drift) and honours the literals extracted from the
reference's own tests.  GPU: the CUDA path, through the C ABI, reproduces the same fixtures WITHOUT the oracle in the
loop (parity mode is bit-exact; tolerances are written where the reference states one).
"""
import json
import os

import numpy as np
import pytest

import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "oracle_vectors.json")) as f:
    VEC = json.load(f)
with open(os.path.join(HERE, "golden", "reference_kats.json"), encoding="utf-8") as f:
    KAT = json.load(f)


def unhex(a, shape=None):
    x = np.array([float.fromhex(s) for s in a])
    return x.reshape(shape) if shape else x


XSQ = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] ** 2), 1, 0)
T_PEND = KAT["pendulum_period_theta_1.3"]["value"]
LAM_REF = [KAT[f"lorenz_lambda{i}_t20"]["value"] for i in (1, 2, 3)]
LAM_TOL = 1e-4  # test/lyapunov.jl:205 mytol


# ---- CPU: oracle against the fixtures --------------------------------------------------------------------------
def test_oracle_reproduces_kepler_vectors():
    import oracle
    k = VEC["kepler28"]
    n = len(k["indices"])
    q = unhex(k["q0"], (4, n))
    q0, mu, _ = P.kepler_ensemble(4096)
    assert np.array_equal(q, q0[:, k["indices"]])
    jets = unhex(k["jets"], k["jets_shape"])
    for j in range(n):
        xj, _, dt = oracle.jet(P.kepler_tape(), 28, q[:, j], mu, abstol=1e-20)
        assert np.array_equal(xj, jets[j]) and dt == float.fromhex(k["dt"][j])
    s = oracle.integrate(P.kepler_tape(), 28, q, 0.0, 2 * np.pi * k["tmax_periods"], mu, maxsteps=2000)
    assert np.array_equal(s.xfinal, unhex(k["xfinal"], (4, n))) and s.nsteps.tolist() == k["nsteps"]


def test_oracle_honours_reference_literals():
    import oracle
    xj, _, dt = oracle.jet(XSQ, 28, [3.0], abstol=1e-20)
    assert KAT["xsq_last_coeff_is_3_pow_order_plus_1"]["value"] and xj[0, -1] == 3.0 ** 29
    assert np.array_equal(xj[0], unhex(VEC["xsq28"]["jet"])) and dt == float.fromhex(VEC["xsq28"]["dt"])
    for tmax, key in ((0.3, "xsq_len_tv_tmax_0.3"), (0.33, "xsq_len_tv_tmax_0.33")):
        s = oracle.integrate(XSQ, 28, [3.0], 0.0, tmax, abstol=1e-20)
        assert int(s.nsteps[0]) + 1 == KAT[key]["value"] == KAT[key.replace("xsq_", "xsq_many_")]["value"]
    pe = VEC["pendulum_events28"]
    assert pe["nevents"] == 5
    assert np.max(np.abs(unhex(pe["tevents"]) - np.arange(1, 6) * T_PEND / 2)) < 1e-13  # test/rootfinding.jl:120-122
    lam = unhex(VEC["lorenz_lyap28"]["lam"])
    assert all(abs(lam[i] - LAM_REF[i]) < LAM_TOL for i in range(3))


# ---- GPU: the CUDA path against the fixtures (no oracle in the loop) ---------------------------------------------
@pytest.mark.gpu
def test_gpu_kepler_jets_steps_and_final_states_match_golden():
    k = VEC["kepler28"]
    n = len(k["indices"])
    q = unhex(k["q0"], (4, n))
    mu = np.array([-1.0])
    kern = tb.get_kernel(P.kepler_tape(), 28, L.KIND_PLAIN, L.MODE_PARITY)
    xj, dt = kern.jet(q, mu, abstol=1e-20)
    jets = unhex(k["jets"], k["jets_shape"])  # [n, dof, p+1]
    assert np.array_equal(np.transpose(xj, (2, 0, 1)), jets)      # bit-exact (north_star allows 2 ulp)
    assert np.array_equal(dt, unhex(k["dt"]))
    s = tb.taylorinteg(P.kepler_tape(), q, 0.0, 2 * np.pi * k["tmax_periods"], 28, 1e-20, mu, maxsteps=2000)
    assert np.array_equal(s.xfinal, unhex(k["xfinal"], (4, n))) and s.nsteps.tolist() == k["nsteps"]


@pytest.mark.gpu
def test_gpu_xsquared_matches_reference_literals():
    kern = tb.get_kernel(XSQ, 28, L.KIND_PLAIN, L.MODE_PARITY)
    xj, dt = kern.jet(np.array([[3.0]]), None, abstol=1e-20)
    assert np.array_equal(xj[0, :, 0], 3.0 ** np.arange(1, 30))    # test/one_ode.jl:203, test/many_ode.jl:164
    assert dt[0] == float.fromhex(VEC["xsq28"]["dt"])
    for tmax, key in ((0.3, "xsq_len_tv_tmax_0.3"), (0.33, "xsq_len_tv_tmax_0.33")):
        s = tb.taylorinteg(XSQ, np.array([3.0]), 0.0, tmax, 28, 1e-20)
        assert s.t.size == KAT[key]["value"] and s.t[-1] == tmax   # test/one_ode.jl:219-221, 239-241
    s = tb.taylorinteg(XSQ, np.array([3.0]), 0.0, 0.33, 28, 1e-20)
    assert np.array_equal(s.t, unhex(VEC["xsq28"]["tv"])) and np.array_equal(s.x[:, 0], unhex(VEC["xsq28"]["xv"]))
    assert np.array_equal(s.p[0, 0], 3.0 ** np.arange(1, 30))       # psol[1] == Taylor1([3.0^i for i = 1:order+1])


@pytest.mark.gpu
def test_gpu_pendulum_events_match_golden():
    pe = VEC["pendulum_events28"]
    s = tb.taylorinteg_events(P.pendulum_tape(), np.array([1.3, 0.0]), 0.0, float.fromhex(pe["tmax"]), 28, 1e-20,
                              maxsteps=1000)
    assert s.tevents.size == pe["nevents"] == 5
    assert np.max(np.abs(s.tevents - np.arange(1, 6) * T_PEND / 2)) < 1e-13          # test/rootfinding.jl:120-122
    assert np.array_equal(s.tevents, unhex(pe["tevents"])) and np.array_equal(s.gresids, unhex(pe["gresids"]))
    assert np.array_equal(s.xevents.T.reshape(-1), unhex(pe["xevents"]))
    assert np.array_equal(s.x[-1], unhex(pe["xfinal"])) and s.t.size - 1 == pe["nsteps"]


@pytest.mark.gpu
def test_gpu_lorenz_lyapunov_matches_golden():
    g = VEC["lorenz_lyap28"]
    s = tb.lyap_taylorinteg(P.lorenz_tape(), np.array(g["q0"]), 0.0, g["tmax"], 28, 1e-20, np.array(g["params"]),
                            maxsteps=2000)
    lam = s.lam[-1]
    assert all(abs(lam[i] - LAM_REF[i]) < LAM_TOL for i in range(3))                 # test/lyapunov.jl:208-210
    assert np.array_equal(lam, unhex(g["lam"])) and np.array_equal(s.x[-1], unhex(g["xfinal"]))
    assert s.t.size - 1 == g["nsteps"]
