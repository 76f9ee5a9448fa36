#!/bin/bash
# quick C2 timing of the library in the working tree (device-resident, B = 4096 and 8192)
cat > /tmp/c2_bench.py <<'PY'
import sys, time, os
sys.path.insert(0, os.getcwd())
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
pb = synth.config2(batch=B, horizon=100)
with Handle(pb) as h:
    for rep in range(5):
        h.upload_u0(None)
        h.synchronize()
        t0 = time.perf_counter()
        h.run_recursive(20, True, False)
        h.synchronize()
        dt = time.perf_counter() - t0
    h.set_profiling(True)
    h.upload_u0(None)
    h.run_recursive(20, True, False)
    h.synchronize()
    pr = h.profile()
    print("B %d ms %.3f solves/s %.0f backward %.3f forward %.3f" % (B, dt * 1e3, pb.batch / dt, pr["riccati_backward"][0], pr["linesearch_rollout"][0]))
PY
for B in 4096 8192; do python /tmp/c2_bench.py $B; done
