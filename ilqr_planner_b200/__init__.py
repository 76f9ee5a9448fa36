"""ilqr_planner_b200 — B200-native batched iLQR solve path (drop-in for idiap/ilqr_planner's Solver::solve).

The product is the CUDA library behind the C ABI in include/ilqrb.h (ilqr_planner_b200/libilqrb.so, sources in
ilqr_planner_b200/csrc/).  This package holds the host-side mirror of the reference interface and plumbing.
"""
from . import cabi  # noqa: F401
