"""debug: config5 (n=m=14) fast path vs oracle, per-quantity relative errors"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
from oracle import oracle
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 6
which = sys.argv[2] if len(sys.argv) > 2 else "c5"
pb = {"c5": synth.config5, "c2": synth.config2, "c1": synth.config1}[which](batch=16)
with Handle(pb) as h:
    out = h.solve_recursive(None, nb, True, False)
ref = oracle.solve_recursive(pb, None, nb, True, False, threads=8)
def rel(a, b, s=None):
    s = s or np.nanmax(np.abs(b)); return float(np.nanmax(np.abs(a - b)) / s)
print("alpha equal", np.array_equal(out.alpha_hist, ref.alpha_hist))
print("margins min", np.nanmin(ref.margin_hist, axis=1))
for nm in ("cost_hist", "X", "U", "K", "d", "fX"):
    a, b = getattr(out, nm), getattr(ref, nm)
    print(nm, "rel", rel(a, b), "max|ref|", np.nanmax(np.abs(b)), "per-problem", np.array2string(np.nanmax(np.abs(a - b).reshape(16, -1), axis=1), precision=2))
