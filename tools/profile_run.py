"""Small fixed workload for ncu: config 2 (B=4096, T=100), a few iterations of the Recursive solver."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
nb_iter = int(sys.argv[1]) if len(sys.argv) > 1 else 3
which = sys.argv[2] if len(sys.argv) > 2 else "c2"
pb = {"c2": lambda: synth.config2(batch=4096), "c3": lambda: synth.config3(batch=4096), "c5": lambda: synth.config5(batch=4096)}[which]()
with Handle(pb) as h:
    h.upload_u0(None)
    h.run_recursive(nb_iter, True, False)
    h.synchronize()
    print("launches", h.launch_count())
