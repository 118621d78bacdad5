"""Random right-hand sides: tracer + tape oracle against tape-free series arithmetic (tests/pygeneric.py).

The reference's strongest invariant is parsed == generic bit for bit (test/taylorize.jl:1267-1273, 1354-1364); here it is
asked of RANDOM expression trees, so that the tape front end (operand kinds, scalar folding, retargeting of results to
dx, aux series, reuse of sub-expressions) is checked on shapes nobody wrote by hand.  CPU only."""
import random

import numpy as np

import oracle
import pygeneric as G
import taylorintegration_b200 as tb
from taylorintegration_b200 import tape as TP

UN = ("sin", "cos", "exp_b", "sqrt_p", "log_p", "pow15_p", "sq", "cube", "neg",
      "tan_b", "tanh", "sinh_b", "cosh_b", "atan", "asin_b", "acos_b")  # the second row: primitives with an aux series
BIN = ("add", "sub", "mul", "div_p", "addc", "mulc", "csub", "cdiv_p")


def gen(rng, dof, npar, depth):
    """Expression tree over x[i], t, p[j] and literals.  `_p` operators get a positive argument (2.5 + sin(e)), `_b` a
    bounded one, so every value stays well inside the domain of its primitive."""
    if depth == 0 or rng.random() < 0.15:
        r = rng.random()
        if r < 0.7:
            return ("x", rng.randrange(dof))
        if r < 0.85:
            return ("t",)
        return ("xp", rng.randrange(dof), rng.randrange(npar)) if npar else ("x", rng.randrange(dof))
    if rng.random() < 0.45:
        return (rng.choice(UN), gen(rng, dof, npar, depth - 1))
    op = rng.choice(BIN)
    c = round(rng.uniform(0.25, 2.0), 3)
    return (op, gen(rng, dof, npar, depth - 1), gen(rng, dof, npar, depth - 1), c)


def ev(n, x, p, t, F):
    k = n[0]
    if k == "x":
        return x[n[1]]
    if k == "t":
        return t
    if k == "xp":
        return p[n[2]] * x[n[1]]
    if k in UN:
        a = ev(n[1], x, p, t, F)
        if k == "sin":
            return F.sin(a)
        if k == "cos":
            return F.cos(a)
        if k == "exp_b":
            return F.exp(0.3 * F.sin(a))
        if k == "sqrt_p":
            return F.sqrt(2.5 + F.sin(a))
        if k == "log_p":
            return F.log(2.5 + F.cos(a))
        if k == "pow15_p":
            return (2.5 + F.sin(a)) ** 1.5
        if k == "sq":
            return a ** 2
        if k == "cube":
            return F.sin(a) ** 3
        if k == "tan_b":
            return F.tan(0.5 * F.sin(a))
        if k == "tanh":
            return F.tanh(F.sin(a))
        if k == "sinh_b":
            return F.sinh(F.cos(a))
        if k == "cosh_b":
            return F.cosh(F.sin(a))
        if k == "atan":
            return F.atan(a)
        if k == "asin_b":
            return F.asin(0.5 * F.sin(a))
        if k == "acos_b":
            return F.acos(0.5 * F.cos(a))
        return -a
    a, b, c = ev(n[1], x, p, t, F), ev(n[2], x, p, t, F), n[3]
    if k == "add":
        return a + b
    if k == "sub":
        return a - b
    if k == "mul":
        return F.sin(a) * F.cos(b)
    if k == "div_p":
        return a / (2.5 + F.sin(b))
    if k == "addc":
        return c + a
    if k == "mulc":
        return a * c
    if k == "csub":
        return c - F.sin(a)
    return c / (2.5 + F.cos(b))


def test_random_right_hand_sides_generic_equals_parsed_bitwise():
    rng = random.Random(20261105)
    order, checked, nops = 10, 0, 0
    for case in range(60):
        dof = rng.choice((1, 2, 3))
        npar = rng.choice((0, 1, 2))
        trees = [gen(rng, dof, npar, rng.choice((2, 3, 4))) for _ in range(dof)]
        # bounded dynamics: every component is damped towards a bounded forcing
        def rhs(F):
            def f(dx, x, p, t):
                for i in range(dof):
                    dx[i] = F.sin(ev(trees[i], x, p, t, F)) - 0.5 * x[i]
            return f
        q0 = np.array([rng.uniform(-1.0, 1.0) for _ in range(dof)])
        par = np.array([rng.uniform(0.5, 1.5) for _ in range(npar)]) if npar else None
        t0 = rng.uniform(-1.0, 1.0)
        tape = tb.taylorize(rhs(TP), dof, npar)
        xp, _, _ = oracle.jet(tape, order, q0, par, t0=t0)
        xg = np.array(G.jetcoeffs(rhs(G), q0, None if par is None else list(par), t0, order))
        assert np.array_equal(xg, xp), (case, trees)
        checked += 1
        nops += len(tape.ops)
    assert checked == 60 and nops > 600  # a few hundred traced primitives went through both evaluators
