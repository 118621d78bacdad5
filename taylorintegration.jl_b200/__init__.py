"""taylor-b200: the ensemble `taylorinteg` hot path of TaylorIntegration.jl on NVIDIA B200 (sm_100a).

The package directory carries the reference's name (`taylorintegration.jl_b200`), which is not a valid Python
identifier; import it through the repo-root shim `taylorintegration_b200`.
"""
from . import _lib  # noqa: F401
from ._lib import TaylorB200Error  # noqa: F401
from .tape import (Tape, taylorize, sin, cos, sqrt, exp, log, zero, tan, sinh, cosh, tanh, atan, asin,  # noqa: F401
                   acos)
from .api import (EnsembleSolution, Kernel, TaylorSolution, fp64_peak, get_kernel, lyap_taylorinteg, make_opts,  # noqa: F401
                  taylorinteg, taylorinteg_events)
from .jettransport import JetSolution, JetTransport, taylorinteg_jet  # noqa: F401
from . import problems  # noqa: F401
from . import sharding  # noqa: F401
from .sharding import lyap_taylorinteg_sharded, taylorinteg_sharded  # noqa: F401

__all__ = ["taylorinteg", "lyap_taylorinteg", "taylorize", "TaylorSolution", "EnsembleSolution", "Kernel", "Tape",
           "sin", "cos", "sqrt", "exp", "log", "zero", "fp64_peak", "problems", "TaylorB200Error",
           "JetTransport", "JetSolution", "taylorinteg_jet", "sharding", "taylorinteg_sharded",
           "lyap_taylorinteg_sharded"]
from . import solution  # noqa: F401,E402
from .solution import DenseEnsemble, flipsign, join, reverse  # noqa: F401,E402
