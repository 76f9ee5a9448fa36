"""PyLQR — the reference's Python module name (pylqr_planner/src/bindings.cpp:48), served by the B200 solver.

`from PyLQR.sim import KDLRobot`, `from PyLQR.solver import ILQRRecursive` ... work as in the reference's tutorials.
The compiled part (`_pylqr`, built by `__graft_entry__.build()` from ilqr_planner_b200/csrc/pylqr_bindings.cpp) binds the
C++ host API of include/ilqr_planner_b200/; it links against ilqr_planner_b200/libilqrb.so and fails loudly if that
library (the CUDA path) is missing.
"""
import sys as _sys

from . import _pylqr

sim = _pylqr.sim
system = _pylqr.system
solver = _pylqr.solver
utils = _pylqr.utils
for _name, _mod in (("sim", sim), ("system", system), ("solver", solver), ("utils", utils), ("utils.primitives", utils.primitives)):
    _sys.modules[__name__ + "." + _name] = _mod
