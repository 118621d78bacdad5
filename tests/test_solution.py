"""Dense-output consumers: reverse / flipsign / join (src/integrator/taylorsolution.jl:475-511), the serialization
layouts (513-657; test/solution.jl round trip) and the device-resident `sol(t)`."""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P, solution as S


def _oracle_solution(tape, order, q0, t0, tmax, params=None, maxsteps=500):
    o = oracle.integrate(tape, order, q0, t0, tmax, params, steps=True, dense=True, maxsteps=maxsteps)
    n = int(o.nsteps[0])
    return tb.TaylorSolution(t=o.t[: n + 1, 0].copy(), x=o.x[:, : n + 1, 0].T.copy(),
                             p=np.transpose(o.p[:, :, :n, 0], (2, 0, 1)).copy())


def _horner(p, dt):
    s = p[..., -1].copy()
    for k in range(p.shape[-1] - 2, -1, -1):
        s = s * dt + p[..., k]
    return s


def test_serialization_layout_and_round_trip(tmp_path):
    sol = _oracle_solution(P.kepler_tape(), 28, np.array([0.2, 0.0, 0.0, 3.0]), 0.0, 3.0, np.array([-1.0]))
    d = S.to_serialization(sol)
    ns, dof, k = sol.p.shape
    assert int(d["order"]) == 28 and tuple(d["dims"]) == (ns, dof) and d["p"].shape == (ns * dof * k,)
    # p[(i-1)*k+1 : i*k] = coefficients of the i-th polynomial in eachindex order of the (steps x dof) matrix: the
    # matrix is walked column-major, so i-1 = c*ns + s (taylorsolution.jl:540-543)
    for c, s in ((0, 0), (2, 5), (3, ns - 1)):
        i = c * ns + s
        assert np.array_equal(d["p"][i * k:(i + 1) * k], sol.p[s, c])
    back = S.from_serialization(d)
    assert np.array_equal(back.t, sol.t) and np.array_equal(back.p, sol.p) and np.array_equal(back.x, sol.x)
    S.save(str(tmp_path / "sol.npz"), sol)
    again = S.load(str(tmp_path / "sol.npz"))
    assert np.array_equal(again.p, sol.p) and np.array_equal(again.t, sol.t)
    # TaylorN layout: polynomial, time coefficient, homogeneous degree ascending, monomial (taylorsolution.jl:599-609)
    mon = oracle.tn_monomials(2, 3)
    pn = np.arange(3 * 2 * 4 * len(mon), dtype=float).reshape(3, 2, 4, len(mon))
    dn = S.to_serialization_n(np.arange(4.0), pn, mon)
    assert int(dn["varorder"]) == 3 and dn["p"].shape == (pn.size,)
    assert np.array_equal(dn["p"][:len(mon)], pn[0, 0, 0]) and np.array_equal(dn["p"][len(mon) * 4:len(mon) * 5], pn[1, 0, 0])


def test_reverse_flipsign_join():
    tape, q0, mu = P.kepler_tape(), np.array([0.2, 0.0, 0.0, 3.0]), np.array([-1.0])
    fwd = _oracle_solution(tape, 28, q0, 0.0, 2.0, mu)
    bwd = _oracle_solution(tape, 28, q0, 0.0, -2.0, mu)
    r = S.reverse(fwd)
    n = fwd.p.shape[0]
    assert np.array_equal(r.t, fwd.t[::-1]) and r.p.shape == fwd.p.shape
    for i in range(1, n):
        assert np.array_equal(r.p[i], fwd.p[n - i])      # prev[i] = p[end-i+2] (1-based)
    # the first polynomial is the last step's expansion re-centred at the final time: same function
    dtl = fwd.t[-1] - fwd.t[-2]
    for tau in (-0.01, 0.0, 0.02):
        assert np.allclose(_horner(r.p[0], tau), _horner(fwd.p[-1], dtl + tau), rtol=1e-13, atol=1e-14)
    assert np.allclose(r.x[0], fwd.x[-1], rtol=1e-14) and np.allclose(r.x[-1], _horner(r.p[-1], r.t[-1] - r.t[-2]))
    f = S.flipsign(fwd)
    assert np.array_equal(f.t, -fwd.t) and np.array_equal(f.p[..., 1::2], -fwd.p[..., 1::2]) and np.array_equal(f.p[..., ::2], fwd.p[..., ::2])
    j = S.join(bwd, fwd)
    assert j.t[0] == -2.0 and j.t[-1] == 2.0 and np.all(np.diff(j.t) > 0) and j.p.shape[0] == bwd.p.shape[0] + fwd.p.shape[0]
    # the joined solution interpolates both branches: compare with fresh integrations to the same times
    for tq in (-1.3, -0.2, 0.7, 1.9):
        idx = min(np.searchsorted(j.t, tq, side="right") - 1, len(j.t) - 2)
        got = _horner(j.p[idx], tq - j.t[idx])
        ref = oracle.integrate(tape, 28, q0, 0.0, tq, mu).xfinal[:, 0]
        assert np.allclose(got, ref, rtol=1e-9, atol=1e-10)
    with pytest.raises(AssertionError):
        S.join(fwd, _oracle_solution(tape, 28, q0, 0.5, 1.0, mu))


@pytest.mark.gpu
def test_device_resident_dense_eval_and_serialize():
    import torch
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(130)
    tmax = 2 * np.pi
    g = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=100, output="steps")
    de = tb.DenseEnsemble(g, 28)
    tq = np.linspace(0.0, tmax, 17)
    xq = de.eval(tq).cpu().numpy()                     # [dof, nq, N]
    for i in (0, 7, 129):
        s = g.trajectory(i)
        for j, t in enumerate(tq):
            idx = len(s.t) - 2 if t == s.t[-1] else np.searchsorted(s.t, t, side="right") - 1
            ref = _horner(s.p[idx], t - s.t[idx])     # multiply then add, like the reference's evaluate
            assert np.array_equal(xq[:, j, i], ref)
        assert np.array_equal(s(tq), xq[:, :, i].T)    # the host-buffer path gives the same bits
    flat, offs = de.serialize()
    flat, offs = flat.cpu().numpy(), offs.cpu().numpy()
    for i in (0, 64, 129):
        d = S.to_serialization(g.trajectory(i))
        assert np.array_equal(flat[offs[i]:offs[i + 1]], d["p"])
    assert offs[-1] == int(g.nsteps.sum()) * 4 * 29
