"""`taylorize`: trace an ODE right-hand side into a tape of Taylor primitives.

This is the host-side counterpart of the reference's `@taylorize` macro
(src/parse_eqs.jl:1468-1475).  The macro unfolds every assignment of the RHS
body into single unary/binary calls (Espresso `to_expr(ExGraph(simplify(ex)))`,
parse_eqs.jl:698-699) and replaces each call by the in-place TaylorSeries
primitive `TaylorSeries.<op>!(res, a, b, ord)` (parse_eqs.jl:963-1135).  Here
the same unfolding is obtained by operator overloading: the user function
`f(dq, q, p, t)` is called once on symbolic operands and every arithmetic
operation appends one `tb_op_t` (include/taylorb200.h) to the tape, in program
order, exactly one primitive per call.

Also here:
  * the event function `g(dx, x, p, t) -> (Bool, series)` (src/rootfinding.jl:124-132),
  * the Jacobian for Lyapunov spectra: the reference evaluates the RHS on
    `TaylorN{Taylor1}` of order 1 (`stabilitymatrix!`, src/lyapunovspectrum.jl:45-64);
    here the same first-order forward mode is applied to the tape itself and
    yields, per entry, a structural zero, a constant, a parameter multiple or a
    series computed by a third tape segment.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

from . import _lib as L


@dataclass
class Op:
    op: int
    res: int
    a: int = 0
    b: int = 0
    aux: int = 0
    imm: float = 0.0

    def describe(self) -> str:
        def nm(o: int) -> str:
            k, i = L.operand_kind(o), L.operand_index(o)
            return {L.K_NONE: "-", L.K_X: f"x[{i}]", L.K_DX: f"dx[{i}]", L.K_TMP: f"tmp{i}", L.K_PAR: f"p[{i}]",
                    L.K_TIME: "t", L.K_IMM: repr(self.imm)}[k]
        s = f"{nm(self.res)} = {L.OP_NAMES.get(self.op, self.op)}({nm(self.a)}"
        if L.operand_kind(self.b) != L.K_NONE:
            s += f", {nm(self.b)}"
        if self.op == L.OP_POW:
            s += f", {self.imm!r}"
        if L.operand_kind(self.aux) != L.K_NONE:
            s += f"; aux={nm(self.aux)}"
        return s + ")"


class _Tracer:
    def __init__(self, dof: int, npar: int):
        self.dof, self.npar = dof, npar
        self.segments: List[List[Op]] = [[], [], []]  # rhs, event, jacobian
        self.cur = 0
        self.ntmp = 0

    @property
    def ops(self) -> List[Op]:
        return self.segments[self.cur]

    def new_tmp(self) -> int:
        self.ntmp += 1
        return L.operand(L.K_TMP, self.ntmp - 1)

    def emit(self, op: int, a: "Val", b: Optional["Val"] = None, imm: Optional[float] = None, aux: bool = False):
        o = Op(op, self.new_tmp())
        o.a = a.operand
        if a.kind == L.K_IMM:
            o.imm = a.value
        if b is not None:
            o.b = b.operand
            if b.kind == L.K_IMM:
                if a.kind == L.K_IMM:
                    raise ValueError("an op can carry only one literal")
                o.imm = b.value
        if imm is not None:
            o.imm = float(imm)
        if aux:
            o.aux = self.new_tmp()
        self.ops.append(o)
        return o


class Val:
    """A traced operand: a Taylor series (state, derivative, temporary, time) or a number (parameter, literal)."""
    __slots__ = ("tr", "kind", "index", "value")
    __array_priority__ = 1000

    def __init__(self, tr: _Tracer, kind: int, index: int = 0, value: float = 0.0):
        self.tr, self.kind, self.index, self.value = tr, kind, index, value

    @property
    def operand(self) -> int:
        return L.operand(self.kind, self.index)

    @property
    def is_number(self) -> bool:
        return self.kind in (L.K_PAR, L.K_IMM)

    @staticmethod
    def from_operand(tr: _Tracer, o: int) -> "Val":
        return Val(tr, L.operand_kind(o), L.operand_index(o))

    def _lift(self, other) -> "Val":
        if isinstance(other, Val):
            return other
        if isinstance(other, _ParamVector):
            return other._scalar()
        if isinstance(other, (int, float)):
            return Val(self.tr, L.K_IMM, 0, float(other))
        return NotImplemented

    def _bin(self, op: int, other, swap: bool = False) -> "Val":
        o = self._lift(other)
        if o is NotImplemented:
            return NotImplemented
        a, b = (o, self) if swap else (self, o)
        if a.kind == L.K_IMM and b.kind == L.K_IMM:
            f = {L.OP_ADD: lambda x, y: x + y, L.OP_SUB: lambda x, y: x - y, L.OP_MUL: lambda x, y: x * y,
                 L.OP_DIV: lambda x, y: x / y}[op]
            return Val(self.tr, L.K_IMM, 0, f(a.value, b.value))
        return Val.from_operand(self.tr, self.tr.emit(op, a, b).res)

    def __add__(self, o): return self._bin(L.OP_ADD, o)
    def __radd__(self, o): return self._bin(L.OP_ADD, o, True)
    def __sub__(self, o): return self._bin(L.OP_SUB, o)
    def __rsub__(self, o): return self._bin(L.OP_SUB, o, True)
    def __mul__(self, o): return self._bin(L.OP_MUL, o)
    def __rmul__(self, o): return self._bin(L.OP_MUL, o, True)
    def __truediv__(self, o): return self._bin(L.OP_DIV, o)
    def __rtruediv__(self, o): return self._bin(L.OP_DIV, o, True)

    def __neg__(self):
        if self.kind == L.K_IMM:
            return Val(self.tr, L.K_IMM, 0, -self.value)
        if self.kind == L.K_PAR:  # -p is p * (-1): a number stays a number
            return self._bin(L.OP_MUL, -1.0)
        return Val.from_operand(self.tr, self.tr.emit(L.OP_NEG, self).res)

    def __pos__(self):
        return self

    def __pow__(self, n):
        if not isinstance(n, (int, float)):
            raise TypeError("pow! needs a constant real exponent (TaylorSeries pow!(res, a, aux, n, k))")
        if self.is_number:
            # `w^2 * x` with a `local` scalar makes the reference's generated method throw
            # (test/taylorize.jl:953-992); reject it here as well.
            raise TypeError("pow! needs a series base; promote the scalar with `+ zero(t)`")
        return Val.from_operand(self.tr, self.tr.emit(L.OP_POW, self, imm=float(n)).res)

    # comparisons build the Bool of an event tuple (only the constant term is inspected)
    def _gate(self, cmp: int, other) -> "Gate":
        if other != 0:
            raise ValueError("event gates compare a series' constant term against 0")
        return Gate(self, cmp)

    def __gt__(self, o): return self._gate(L.GATE_GT0, o)
    def __lt__(self, o): return self._gate(L.GATE_LT0, o)
    def __ge__(self, o): return self._gate(L.GATE_GE0, o)
    def __le__(self, o): return self._gate(L.GATE_LE0, o)


@dataclass
class Gate:
    val: Val
    cmp: int


class _ParamVector:
    """`p` of `f!(dx, x, p, t)`: indexable, iterable, and usable as a scalar when it has one entry (`local mu = p`)."""

    def __init__(self, tr: _Tracer, n: int):
        self.tr, self.n = tr, n

    def __getitem__(self, i: int) -> Val:
        if not 0 <= i < self.n:
            raise IndexError(i)
        return Val(self.tr, L.K_PAR, i)

    def __len__(self):
        return self.n

    def __iter__(self):
        return (self[i] for i in range(self.n))

    def _scalar(self) -> Val:
        if self.n != 1:
            raise TypeError("params used as a scalar but npar != 1")
        return self[0]

    def __mul__(self, o): return self._scalar() * o
    def __rmul__(self, o): return o * self._scalar()
    def __add__(self, o): return self._scalar() + o
    def __radd__(self, o): return o + self._scalar()
    def __sub__(self, o): return self._scalar() - o
    def __rsub__(self, o): return o - self._scalar()
    def __truediv__(self, o): return self._scalar() / o
    def __rtruediv__(self, o): return o / self._scalar()
    def __neg__(self): return -self._scalar()


class _DxVector:
    """`dx` of `f!`: assignment retargets the producing primitive to `dx[i]` (as `dq[3] = div!(t4, r)` in the
    reference's generated code) or emits `identity!`."""

    def __init__(self, tr: _Tracer):
        self.tr = tr
        self.assigned = [False] * tr.dof

    def __len__(self):
        return self.tr.dof

    def __getitem__(self, i: int) -> Val:
        if not self.assigned[i]:
            raise ValueError(f"dx[{i}] read before assignment")
        return Val(self.tr, L.K_DX, i)

    def __setitem__(self, i: int, v) -> None:
        tr = self.tr
        if not 0 <= i < tr.dof:
            raise IndexError(i)
        if isinstance(v, _ParamVector):
            v = v._scalar()
        if isinstance(v, (int, float)):
            v = Val(tr, L.K_IMM, 0, float(v))
        target = L.operand(L.K_DX, i)
        last = tr.ops[-1] if tr.ops else None
        if v.kind == L.K_TMP and last is not None and last.res == v.operand:
            last.res = target
        elif v.kind == L.K_TMP and last is not None and last.aux == v.operand and last.op in (L.OP_SINCOS, L.OP_SINHCOSH):
            last.aux = target
        else:
            o = Op(L.OP_IDENTITY, target, v.operand)
            if v.kind == L.K_IMM:
                o.imm = v.value
            tr.ops.append(o)
        self.assigned[i] = True


# primitives that carry an auxiliary series (src/parse_eqs.jl:1013-1046)
_AUX_OPS = (L.OP_SINCOS, L.OP_TAN, L.OP_SINHCOSH, L.OP_TANH, L.OP_ATAN, L.OP_ASIN, L.OP_ACOS)


def _unary(op: int, fallback: Callable[[float], float], pick_aux: bool = False):
    def fn(x):
        if isinstance(x, _ParamVector):
            x = x._scalar()
        if not isinstance(x, Val):
            return fallback(x)
        if x.kind == L.K_IMM:
            return Val(x.tr, L.K_IMM, 0, fallback(x.value))
        if x.is_number:
            raise TypeError("Taylor primitive needs a series argument; promote the scalar with `+ zero(t)`")
        o = x.tr.emit(op, x, aux=(op in _AUX_OPS))
        return Val.from_operand(x.tr, o.aux if pick_aux else o.res)
    return fn


sin = _unary(L.OP_SINCOS, math.sin)
cos = _unary(L.OP_SINCOS, math.cos, pick_aux=True)
sqrt = _unary(L.OP_SQRT, math.sqrt)
exp = _unary(L.OP_EXP, math.exp)
log = _unary(L.OP_LOG, math.log)
tan = _unary(L.OP_TAN, math.tan)        # tan!(c, a, c2, k), aux = tan(a)^2
sinh = _unary(L.OP_SINHCOSH, math.sinh)  # sinhcosh!(s, c, a, k)
cosh = _unary(L.OP_SINHCOSH, math.cosh, pick_aux=True)
tanh = _unary(L.OP_TANH, math.tanh)
atan = _unary(L.OP_ATAN, math.atan)      # aux = 1 + a^2
asin = _unary(L.OP_ASIN, math.asin)      # aux = sqrt(1 - a^2)
acos = _unary(L.OP_ACOS, math.acos)


def zero(x) -> float:
    """`zero(x)`: used as `c + zero(t)` to promote a number; adding 0.0 literal folds away here."""
    return 0.0


# ---- first-order forward mode over the tape (Jacobian) ----------------------
_Z = ("z",)


class _Jac:
    def __init__(self, tr: _Tracer):
        self.tr = tr
        self.d = {}  # operand -> list of dvals, one per state variable

    def V(self, o: int) -> Val:
        return Val.from_operand(self.tr, o)

    def imm(self, c: float) -> Val:
        return Val(self.tr, L.K_IMM, 0, float(c))

    def par(self, p: int) -> Val:
        return Val(self.tr, L.K_PAR, p)

    def get(self, o: int, j: int):
        k = L.operand_kind(o)
        if k == L.K_X:
            return ("n", 1.0, None) if L.operand_index(o) == j else _Z
        if k in (L.K_DX, L.K_TMP):
            return self.d[o][j]
        return _Z

    def ser(self, v) -> Val:
        """dval -> series operand (materialising numbers)"""
        if v[0] == "s":
            return self.V(v[1])
        _, c, p = v
        if p is None:
            return self.V(self.tr.emit(L.OP_IDENTITY, self.imm(c)).res)
        return self.V(self.tr.emit(L.OP_MUL, self.par(p), self.imm(c)).res)

    def num_operand(self, v) -> Optional[Val]:
        """dval number usable directly as a numeric operand"""
        _, c, p = v
        if p is None:
            return self.imm(c)
        if c == 1.0:
            return self.par(p)
        return None

    def neg(self, v):
        if v[0] == "z":
            return v
        if v[0] == "n":
            return ("n", -v[1], v[2])
        return ("s", self.tr.emit(L.OP_NEG, self.V(v[1])).res)

    def mul(self, v, x: Val):
        """dval * (number or series operand)"""
        if v[0] == "z":
            return v
        if v[0] == "n":
            _, c, p = v
            if x.kind == L.K_IMM:
                return ("n", c * x.value, p) if c * x.value != 0.0 else _Z
            if x.kind == L.K_PAR:
                if p is None:
                    return ("n", c, x.index)
                return ("s", self.tr.emit(L.OP_MUL, self.ser(v), x).res)
            if p is None:
                if c == 1.0:
                    return ("s", x.operand)
                if c == -1.0:
                    return ("s", self.tr.emit(L.OP_NEG, x).res)
                return ("s", self.tr.emit(L.OP_MUL, self.imm(c), x).res)
            t = self.tr.emit(L.OP_MUL, self.par(p), x).res
            return ("s", t) if c == 1.0 else ("s", self.tr.emit(L.OP_MUL, self.imm(c), self.V(t)).res)
        if x.kind == L.K_IMM and x.value == 1.0:
            return v
        return ("s", self.tr.emit(L.OP_MUL, self.V(v[1]), x).res)

    def add(self, u, v, sub: bool = False):
        if v[0] == "z":
            return u
        if u[0] == "z":
            return self.neg(v) if sub else v
        opc = L.OP_SUB if sub else L.OP_ADD
        if u[0] == "n" and v[0] == "n":
            if u[2] == v[2]:
                c = u[1] - v[1] if sub else u[1] + v[1]
                return ("n", c, u[2]) if c != 0.0 else _Z
            return ("s", self.tr.emit(opc, self.ser(u), self.ser(v)).res)
        if u[0] == "s" and v[0] == "n":
            nv = self.num_operand(v)
            b = nv if nv is not None else self.ser(v)
            return ("s", self.tr.emit(opc, self.V(u[1]), b).res)
        if u[0] == "n" and v[0] == "s":
            nu = self.num_operand(u)
            a = nu if nu is not None else self.ser(u)
            return ("s", self.tr.emit(opc, a, self.V(v[1])).res)
        return ("s", self.tr.emit(opc, self.V(u[1]), self.V(v[1])).res)

    def div_series(self, v, s: Val):
        if v[0] == "z":
            return v
        if v[0] == "n":
            nv = self.num_operand(v)
            a = nv if nv is not None else self.ser(v)
            return ("s", self.tr.emit(L.OP_DIV, a, s).res)
        return ("s", self.tr.emit(L.OP_DIV, self.V(v[1]), s).res)

    def div_number(self, v, x: Val):
        if v[0] == "z":
            return v
        if v[0] == "n" and x.kind == L.K_IMM:
            return ("n", v[1] / x.value, v[2])
        return ("s", self.tr.emit(L.OP_DIV, self.ser(v), x).res)

    def run(self, ops: Sequence[Op], dof: int):
        tr = self.tr
        for op in ops:
            a, b, r = op.a, op.b, op.res
            ka, kb = L.operand_kind(a), L.operand_kind(b)
            num_a = ka in (L.K_PAR, L.K_IMM)
            num_b = kb in (L.K_PAR, L.K_IMM)
            Va = Val(tr, ka, L.operand_index(a), op.imm) if ka != L.K_NONE else None
            Vb = Val(tr, kb, L.operand_index(b), op.imm) if kb != L.K_NONE else None
            out, out_aux = [], []
            opc, pw = op.op, op.imm
            if opc == L.OP_POW:
                opc = {0.0: L.OP_ONE, 1.0: L.OP_IDENTITY, 2.0: L.OP_SQR, 0.5: L.OP_SQRT}.get(pw, L.OP_POW)
            quot = None
            for j in range(dof):
                da, db = self.get(a, j), self.get(b, j)
                if opc == L.OP_IDENTITY:
                    v = da
                elif opc == L.OP_NEG:
                    v = self.neg(da)
                elif opc in (L.OP_ZERO, L.OP_ONE):
                    v = _Z
                elif opc == L.OP_ADD:
                    v = self.add(da, db)
                elif opc == L.OP_SUB:
                    v = self.add(da, db, sub=True)
                elif opc == L.OP_MUL:
                    v = self.add(self.mul(da, Vb), self.mul(db, Va))
                elif opc == L.OP_DIV:
                    if num_b:
                        v = self.div_number(da, Vb)
                    else:
                        v = self.div_series(self.add(da, self.mul(db, self.V(r)), sub=True), Vb)
                elif opc == L.OP_SQR:
                    v = self.mul(self.mul(da, self.imm(2.0)), Va)
                elif opc == L.OP_SQRT:
                    v = self.div_series(self.mul(da, self.imm(0.5)), self.V(r))
                elif opc == L.OP_POW:
                    if da[0] == "z":
                        v = _Z
                    else:
                        if quot is None:
                            quot = self.V(tr.emit(L.OP_DIV, self.V(r), Va).res)
                        v = self.mul(self.mul(da, self.imm(pw)), quot)
                elif opc == L.OP_EXP:
                    v = self.mul(da, self.V(r))
                elif opc == L.OP_LOG:
                    v = self.div_series(da, Va)
                elif opc == L.OP_SINCOS:
                    v = self.mul(da, self.V(op.aux))
                    out_aux.append(self.neg(self.mul(da, self.V(r))))
                else:
                    raise NotImplementedError(f"no derivative rule for opcode {opc}")
                out.append(v)
            self.d[r] = out
            if op.op == L.OP_SINCOS:
                self.d[op.aux] = out_aux
            del num_a


@dataclass
class Tape:
    """A validated ODE description ready for `tb_ode_create` (and, in tests, for the CPU oracle)."""
    dof: int
    npar: int
    ntmp: int
    ops: List[Op]
    ev_ops: List[Op] = field(default_factory=list)
    jac_ops: List[Op] = field(default_factory=list)
    ev_g: int = 0
    ev_gate: int = 0
    ev_gate_cmp: int = L.GATE_ALWAYS
    jac: Optional[List[Tuple[int, int, float]]] = None  # row-major dof x dof of (kind, operand, c)
    name: str = "ode"

    def all_ops(self) -> List[Op]:
        return list(self.ops) + list(self.ev_ops) + list(self.jac_ops)

    def c_desc(self):
        """(TbOdeDesc, keepalive) for the C ABI."""
        ops = self.all_ops()
        arr = (L.TbOp * len(ops))()
        for i, o in enumerate(ops):
            arr[i] = L.TbOp(o.op, o.res, o.a, o.b, o.aux, 0, o.imm)
        d = L.TbOdeDesc()
        d.dof, d.npar, d.ntmp = self.dof, self.npar, self.ntmp
        d.nops, d.ev_nops, d.jac_nops = len(self.ops), len(self.ev_ops), len(self.jac_ops)
        d.ops = C.cast(arr, C.POINTER(L.TbOp))
        d.ev_g, d.ev_gate, d.ev_gate_cmp = self.ev_g, self.ev_gate, self.ev_gate_cmp
        keep = [arr]
        if self.jac is not None:
            jarr = (L.TbJacEntry * (self.dof * self.dof))()
            for i, (k, o, c) in enumerate(self.jac):
                jarr[i] = L.TbJacEntry(k, o, c)
            d.has_jac = 1
            d.jac = C.cast(jarr, C.POINTER(L.TbJacEntry))
            keep.append(jarr)
        return d, keep

    def describe(self) -> str:
        out = [f"# {self.name}: dof={self.dof} npar={self.npar} ntmp={self.ntmp}"]
        out += [o.describe() for o in self.ops]
        if self.ev_ops or self.ev_g:
            out.append("# event")
            out += [o.describe() for o in self.ev_ops]
        if self.jac_ops:
            out.append("# jacobian")
            out += [o.describe() for o in self.jac_ops]
        return "\n".join(out)


class _JacMatrix:
    """`jac` of a user `jacobianfunc!(jac, x, p, t)` (src/lyapunovspectrum.jl:65-78): `jac[i, j] = value` records the
    entry; a number stays a number (the reference's `-σ + zero(x[1])` constant series multiplies with the same bits)."""

    def __init__(self, tr: _Tracer, dof: int):
        self.tr, self.dof = tr, dof
        self.ent = {}

    def __setitem__(self, ij, v) -> None:
        i, j = ij
        if not (0 <= i < self.dof and 0 <= j < self.dof):
            raise IndexError(ij)
        if isinstance(v, _ParamVector):
            v = v._scalar()
        if isinstance(v, (int, float)):
            self.ent[(i, j)] = (L.JAC_ZERO, 0, 0.0) if v == 0 else (L.JAC_CONST, 0, float(v))
        elif isinstance(v, Val):
            if v.kind == L.K_IMM:
                self.ent[(i, j)] = (L.JAC_ZERO, 0, 0.0) if v.value == 0 else (L.JAC_CONST, 0, float(v.value))
            elif v.kind == L.K_PAR:
                self.ent[(i, j)] = (L.JAC_PAR, v.operand, 1.0)
            else:
                self.ent[(i, j)] = (L.JAC_SERIES, v.operand, 0.0)
        else:
            raise TypeError("a Jacobian entry is a number, a parameter or a series")

    def entries(self):
        missing = [(i, j) for i in range(self.dof) for j in range(self.dof) if (i, j) not in self.ent]
        if missing:
            raise ValueError(f"the Jacobian function does not assign jac{missing}")
        return [self.ent[(i, j)] for i in range(self.dof) for j in range(self.dof)]


def taylorize(f: Callable, dof: int, npar: int = 0, *, event: Optional[Callable] = None, jacobian=False,
              name: Optional[str] = None) -> Tape:
    """Trace `f(dx, x, p, t)` (and optionally an event function `g(dx, x, p, t) -> (gate, series)`) into a Tape.

    `jacobian=True` adds the variational-equation Jacobian needed by `lyap_taylorinteg` (first-order forward mode on
    the tape: the reference's TaylorN path, src/lyapunovspectrum.jl:45-63); `jacobian=jf` traces the user's
    `jf(jac, x, p, t)` instead (the `jacobianfunc!` path, src/lyapunovspectrum.jl:65-78).
    """
    tr = _Tracer(dof, npar)
    x = [Val(tr, L.K_X, i) for i in range(dof)]
    dx = _DxVector(tr)
    p = _ParamVector(tr, npar)
    t = Val(tr, L.K_TIME, 0)
    f(dx, x, p, t)
    missing = [i for i, a in enumerate(dx.assigned) if not a]
    if missing:
        raise ValueError(f"the RHS does not assign dx{missing}")
    tape = Tape(dof, npar, 0, tr.segments[0], name=name or getattr(f, "__name__", "ode"))
    if event is not None:
        tr.cur = 1
        res = event(dx, x, p, t)
        gate, g = res
        if isinstance(g, (int, float)) or (isinstance(g, Val) and g.is_number):
            raise TypeError("the event function must return a series")
        tape.ev_ops = tr.segments[1]
        tape.ev_g = g.operand
        if isinstance(gate, Gate):
            tape.ev_gate, tape.ev_gate_cmp = gate.val.operand, gate.cmp
        elif gate is True:
            tape.ev_gate_cmp = L.GATE_ALWAYS
        else:
            raise TypeError("the event gate must be `True` or a comparison of a series with 0")
    if callable(jacobian):
        tr.cur = 2
        Jm = _JacMatrix(tr, dof)
        jacobian(Jm, x, p, t)
        tape.jac = Jm.entries()
        tape.jac_ops = tr.segments[2]
    elif jacobian:
        tr.cur = 2
        J = _Jac(tr)
        J.run(tr.segments[0], dof)
        ent = []
        for i in range(dof):
            for j in range(dof):
                v = J.get(L.operand(L.K_DX, i), j)
                if v[0] == "z":
                    ent.append((L.JAC_ZERO, 0, 0.0))
                elif v[0] == "n":
                    ent.append((L.JAC_CONST, 0, v[1]) if v[2] is None else (L.JAC_PAR, L.operand(L.K_PAR, v[2]), v[1]))
                else:
                    ent.append((L.JAC_SERIES, v[1], 0.0))
        tape.jac = ent
        tape.jac_ops = tr.segments[2]
    tape.ntmp = tr.ntmp
    return tape


def lyap_column_tape(tape: Tape) -> Optional[Tape]:
    """The column system of `lyap_jetcoeffs!` (src/lyapunovspectrum.jl:161-211) as a plain tape.

    State `[x; v]` with `x' = f(x)` (the tape itself) and `v' = J(x) v` for ONE column `v` of the fundamental matrix:
    `dv[i] = ((J[i,0]*v[0]) + J[i,1]*v[1]) + ...` in the reference's accumulation order (i3 ascending), every product a
    full `mul!` with the Jacobian series first.  Constant / parameter entries multiply as numbers (the coefficients the
    reference adds for their higher orders are exact zeros), structurally zero entries are skipped.  A team of `dof`
    GPU lanes, one per column, integrates a trajectory with this tape (TB_KIND_LYAPCOL).

    Returns None when an entry cannot be expressed with the same bits (a parameter scaled by c != +-1) or the system
    is too large for the register-tiled generator; the caller then uses the single-thread TB_KIND_LYAP kernel.
    """
    if tape.jac is None:
        return None
    n = tape.dof
    if n > 8 or n < 1:
        return None
    ops = [Op(o.op, o.res, o.a, o.b, o.aux, o.imm) for o in list(tape.ops) + list(tape.jac_ops)]
    ntmp = tape.ntmp

    def tmp() -> int:
        nonlocal ntmp
        ntmp += 1
        return L.operand(L.K_TMP, ntmp - 1)

    imm_operand = L.operand(L.K_IMM, 0)
    negated = {o.res: o.a for o in ops if o.op == L.OP_NEG and L.operand_kind(o.a) in (L.K_X, L.K_DX, L.K_TMP)}
    for i in range(n):
        acc = None  # operand holding the running sum
        for m in range(n):
            kind, operand, c = tape.jac[i * n + m]
            v = L.operand(L.K_X, n + m)
            neg = False
            if kind == L.JAC_ZERO:
                continue
            if kind == L.JAC_CONST:
                term = Op(L.OP_MUL, tmp(), imm_operand, v, 0, float(c))
            elif kind == L.JAC_PAR:
                if c not in (1.0, -1.0):
                    return None
                neg = c == -1.0  # (-p)*v and -(p*v) have the same bits
                term = Op(L.OP_MUL, tmp(), operand, v)
            elif kind == L.JAC_SERIES:
                if operand in negated:
                    # J = -s: mul!(-s, v) and -mul!(s, v) have the same bits (every product and partial sum changes
                    # sign only), and s is usually a series the trajectory part already keeps: one stored series less
                    operand, neg = negated[operand], True
                term = Op(L.OP_MUL, tmp(), operand, v)
            else:
                return None
            ops.append(term)
            t = term.res
            if acc is None:
                if neg:
                    o = Op(L.OP_NEG, tmp(), t)
                    ops.append(o)
                    t = o.res
                acc = t
            else:
                o = Op(L.OP_SUB if neg else L.OP_ADD, tmp(), acc, t)
                ops.append(o)
                acc = o.res
        dxi = L.operand(L.K_DX, n + i)
        if acc is None:
            ops.append(Op(L.OP_ZERO, dxi, L.operand(L.K_X, n + i)))
        else:
            ops.append(Op(L.OP_IDENTITY, dxi, acc))
    # drop entries nothing reads any more (e.g. the negated Jacobian series replaced above)
    while True:
        used = set()
        for o in ops:
            used.update((o.a, o.b))
        keep = [o for o in ops if L.operand_kind(o.res) == L.K_DX or o.res in used or (o.aux and o.aux in used)]
        if len(keep) == len(ops):
            break
        ops = keep
    return Tape(2 * n, tape.npar, ntmp, ops, name=tape.name + "+column")
