"""Compare the fast and generic device paths with the oracle on one config (run twice: with and without ILQRB_GENERIC=1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
from oracle import oracle
cfg, B, nit = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
pb = getattr(synth, cfg)(batch=B)
with Handle(pb) as h:
    out = h.solve_recursive(None, nit, True, False)
ref = oracle.solve_recursive(pb, None, nit, True, False, threads=8)
def rel(a, b): return float(np.nanmax(np.abs(a - b)) / np.nanmax(np.abs(b)))
same = [np.array_equal(out.alpha_hist[b], ref.alpha_hist[b], equal_nan=True) for b in range(B)]
print("generic" if os.environ.get("ILQRB_GENERIC") else "fast", cfg, "alpha-equal", sum(same), "of", B,
      {k: "%.2e" % rel(getattr(out, k)[same], getattr(ref, k)[same]) for k in ("cost_hist", "X", "U", "K", "d", "fX")})
