"""C2 device-resident throughput against the batch size (how far B = 4096 is from filling one B200)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
for B in [int(a) for a in sys.argv[1:]] or [1024, 2048, 4096, 8192, 16384]:
    pb = synth.config2(batch=B)
    with Handle(pb) as h:
        h.upload_u0(None)
        h.run_recursive(20, True, False); h.synchronize()
        t0 = time.perf_counter()
        for _ in range(3): h.run_recursive(20, True, False)
        h.synchronize()
        dt = (time.perf_counter() - t0) / 3
        h.set_profiling(True); h.reset_profile(); h.run_recursive(20, True, False); prof = h.profile(); h.set_profiling(False)
    print("B=%6d  %8.2f ms  %9.0f solves/s   %s" % (B, 1e3 * dt, B / dt, {k: round(v[0], 2) for k, v in prof.items()}), flush=True)
