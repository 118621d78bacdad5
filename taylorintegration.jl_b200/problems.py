"""The ODEs and synthetic ensembles of the five benchmark configurations (BASELINE.json `configs`,
SURVEY.md section 8d), written as the reference writes them.

Initial-condition generators use numpy's PCG64 with fixed seeds (not Julia's RNG): the same arrays are fed to the
GPU kernels, the CPU oracle and the timed CPU baseline.
"""
from __future__ import annotations

import numpy as np

from .tape import cos, sin, sqrt, taylorize  # noqa: F401


# -- config 1: planar Kepler problem, benchmark/kepler_benchmarks.jl:9-19 ---------------------------------------
def kepler1(dq, q, p, t):
    mu = p  # `local μ = p`
    r_p3d2 = (q[0] ** 2 + q[1] ** 2) ** 1.5
    dq[0] = q[2]
    dq[1] = q[3]
    dq[2] = mu * q[0] / r_p3d2
    dq[3] = mu * q[1] / r_p3d2


def kepler_tape():
    return taylorize(kepler1, 4, 1, name="kepler1!")


def kepler_ensemble(N: int, emax: float = 0.9):
    """q0(e) = [1-e, 0, 0, sqrt(1-e^2)/(1-e)] (docs/src/kepler.md:63-69 with a = mu = m = 1; e = 0.8 gives the
    benchmark's [0.2, 0, 0, 3.0], kepler_benchmarks.jl:81), e_i = emax*i/(N-1); params mu = -1."""
    e = emax * np.arange(N, dtype=np.float64) / max(N - 1, 1)
    q0 = np.zeros((4, N))
    q0[0] = 1.0 - e
    q0[3] = np.sqrt(1.0 - e * e) / (1.0 - e)
    return q0, np.array([-1.0]), e


KEPLER = dict(order=28, abstol=1e-20, t0=0.0, tmax=2 * np.pi * 100, maxsteps=10_000)


# -- config 2: simple pendulum with events, test/rootfinding.jl:20-26 -------------------------------------------
def pendulum(dx, x, p, t):
    dx[0] = x[1]
    dx[1] = -sin(x[0])


def pendulum_event(dx, x, p, t):
    return (True, x[1])


def pendulum_tape(events: bool = True):
    return taylorize(pendulum, 2, 0, event=pendulum_event if events else None, name="pendulum!")


def pendulum_ensemble(N: int, seed: int = 20261104):
    rng = np.random.Generator(np.random.PCG64(seed))
    th = rng.uniform(0.1, 3.1, size=N)
    if N > 0:
        th[0] = 1.3  # golden period 7.019250311844546 (test/rootfinding.jl:30)
    q0 = np.zeros((2, N))
    q0[0] = th
    return q0


PENDULUM = dict(order=25, abstol=1e-20, t0=0.0, tmax=100.0, maxsteps=5_000, event_capacity=64)


# -- config 3: Lorenz system Lyapunov spectrum, benchmark/Lyap_benchmark.jl:3-9 ---------------------------------
def lorenz(dx, x, p, t):
    sigma, rho, beta = p
    dx[0] = sigma * (x[1] - x[0])
    dx[1] = x[0] * (rho - x[2]) - x[1]
    dx[2] = x[0] * x[1] - beta * x[2]


def lorenz_tape(jacobian: bool = True):
    return taylorize(lorenz, 3, 3, jacobian=jacobian, name="lorenz!")


def lorenz_ensemble(N: int):
    """sigma = 16, beta = 4, rho_i = 28 + 64*i/(N-1) (contains the reference's 45.92 neighbourhood); q0 = [19,20,50]."""
    q0 = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, N))
    par = np.empty((3, N))
    par[0] = 16.0
    par[1] = 28.0 + 64.0 * np.arange(N) / max(N - 1, 1)
    par[2] = 4.0
    return q0, par


LORENZ = dict(order=28, abstol=1e-20, t0=0.0, tmax=100.0, maxsteps=20_000)


# -- config 4: classical many-spin system, benchmark/manyspin_benchmarks.jl:10-96 --------------------------------
def fix_J(nn: int, alpha: float = 1.4) -> np.ndarray:
    """Long-range couplings J[i, i+d] = J[i+d, i] = 1/d^alpha (manyspin_benchmarks.jl:72-82)."""
    J = np.zeros((nn, nn))
    for d in range(1, nn):
        val = 1.0 / d ** alpha
        for i in range(nn - d):
            J[i, i + d] = val
            J[i + d, i] = val
    return J


def fix_B(nn: int, seed: int = 20261104) -> np.ndarray:
    """Local fields.  The reference tabulates 20 values in [0.8, 1.2) (manyspin_benchmarks.jl:46-69) and asserts
    nn < 21; for larger systems the same range is drawn from a seeded generator."""
    return np.random.Generator(np.random.PCG64(seed)).uniform(0.8, 1.2, size=nn)


def manyspin_tape(nn: int, B: np.ndarray = None, J: np.ndarray = None):
    """`spin_odes!` (manyspin_benchmarks.jl:10-41) with `nn`, `B`, `J` as the `local` numbers they are there."""
    B = fix_B(nn) if B is None else np.asarray(B, dtype=np.float64)
    J = fix_J(nn) if J is None else np.asarray(J, dtype=np.float64)

    def spin_odes(dq, q, p, t):
        x = [q[3 * i] for i in range(nn)]
        y = [q[3 * i + 1] for i in range(nn)]
        z = [q[3 * i + 2] for i in range(nn)]
        for i in range(nn):
            sum_s = 0.0  # zero(x[1])
            for a in range(nn):
                xJ = float(J[a, i]) * x[a]
                sum_s = sum_s + xJ
            dq[3 * i] = float(-B[i]) * y[i]
            dq[3 * i + 1] = float(B[i]) * x[i] + z[i] * sum_s
            dq[3 * i + 2] = (-y[i]) * sum_s

    return taylorize(spin_odes, 3 * nn, 0, name=f"spin_odes!({nn})")


def manyspin_ensemble(nn: int, N: int, seed: int = 20261104) -> np.ndarray:
    """Random spins on the unit sphere as in `ini_conditions!` (manyspin_benchmarks.jl:83-96); q0 [3 nn, N]."""
    rng = np.random.Generator(np.random.PCG64(seed + 1))
    sz = 2.0 * (rng.random((nn, N)) - 0.5)
    pha = 2.0 * np.pi * rng.random((nn, N))
    a = np.sqrt(1.0 - sz * sz)
    q0 = np.empty((3 * nn, N))
    q0[0::3], q0[1::3], q0[2::3] = a * np.cos(pha), a * np.sin(pha), sz
    return q0


MANYSPIN = dict(order=25, abstol=1e-20, t0=0.0, tmax=10.0, maxsteps=5000, nspin=100)
