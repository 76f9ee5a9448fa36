"""Small fixed workload for ncu: a few iterations of the Recursive solver (c2, c3, c5, time = PosOrnTime T=100) or of AL-iLQR (c3al)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle, make_al_desc
nb_iter = int(sys.argv[1]) if len(sys.argv) > 1 else 3
which = sys.argv[2] if len(sys.argv) > 2 else "c2"
batch = int(sys.argv[3]) if len(sys.argv) > 3 else (16384 if which == "c3al" else 4096)
if which == "time":
    pb = synth.config4(batch=batch, horizon=100)
else:
    pb = {"c2": synth.config2, "c3": synth.config3, "c3al": synth.config3, "c5": synth.config5}[which](batch=batch)
with Handle(pb) as h:
    h.upload_u0(synth.u0_time(pb) if which == "time" else None)
    if which == "c3al":
        h.run_al(make_al_desc(**synth.al_config3()), nb_iter, True, False)
    else:
        h.run_recursive(nb_iter, True, False)
    h.synchronize()
    print("launches", h.launch_count())
