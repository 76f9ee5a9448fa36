#!/bin/bash
# same-box A/B of the working tree against the build in _ab/ (a worktree of an earlier commit): C2 device-resident solve
cat > /tmp/c2_bench.py <<'PY'
import sys, time, os
sys.path.insert(0, os.getcwd())
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
pb = synth.config2(batch=B, horizon=100)
with Handle(pb) as h:
    for rep in range(4):
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
for i in 1 2; do
  echo new; python /tmp/c2_bench.py 4096
  echo old; (cd _ab && python /tmp/c2_bench.py 4096)
done
sed -e 's/synth.config2(batch=B, horizon=100)/synth.config5(batch=B)/' /tmp/c2_bench.py > /tmp/c5_bench.py
echo "c5 new"; python /tmp/c5_bench.py 8192
echo "c5 new groups=2"; ILQRB_GROUPS=2 python /tmp/c5_bench.py 8192
echo "c5 old"; (cd _ab && python /tmp/c5_bench.py 8192)
