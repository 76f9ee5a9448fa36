"""Device-resident throughput of the other BASELINE.json configs (parity-test shapes, not the bench.py line):
C1 Robot2D, C3 double integrator Recursive and AL-iLQR, C5 dual arm, time-system Recursive."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle, make_al_desc

def run(name, pb, nit, al=None, u0=None, reps=3):
    with Handle(pb) as h:
        h.upload_u0(u0)
        ald = make_al_desc(**al) if al else None
        def go():
            if ald is not None: h.run_al(ald, nit, True, False)
            else: h.run_recursive(nit, True, False)
        go(); h.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps): go()
        h.synchronize()
        dt = (time.perf_counter() - t0) / reps
        h.set_profiling(True); h.reset_profile(); go(); prof = h.profile(); h.set_profiling(False)
    print("%-34s B=%6d it=%3d  %8.2f ms/solve-batch  %10.0f solves/s  %7.3f ms/iteration   %s" % (
        name, pb.batch, nit, 1e3 * dt, pb.batch / dt, 1e3 * dt / nit, {k: round(v[0], 2) for k, v in prof.items()}))

which = sys.argv[1:] or ["c1", "c3", "c3al", "c5", "time"]
if "c1" in which: run("C1 Robot2D position T=50", synth.config1(batch=4096), 20)
if "c3" in which: run("C3 shape Recursive nd=2 T=100", synth.config3(batch=4096), 20)
if "c3al" in which: run("C3 AL-iLQR nd=2 T=100", synth.config3(batch=16384), 40, al=synth.al_config3())
if "c5" in which: run("C5 dual arm n=m=14 T=100", synth.config5(batch=8192), 20)
if "time" in which:
    pb = synth.config4(batch=4096, horizon=100)
    run("PosOrnTime Recursive T=100", pb, 20, u0=synth.u0_time(pb))
