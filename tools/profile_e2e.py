"""Small fixed workload for ncu: one blocking host-buffer solve of C2 (H2D + solve + pack + D2H), to list the kernels behind the downloads."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
from ilqr_planner_b200.problem import ResultBuffers
pb = synth.config2(batch=4096, horizon=100)
out = ResultBuffers(pb, 2, "recursive", pinned=True)
with Handle(pb) as h:
    h.solve_recursive(None, 2, True, False, out=out)
    h.synchronize()
