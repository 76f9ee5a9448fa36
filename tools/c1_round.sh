#!/bin/bash
# column-per-lane backward kernel for n < 7: parity suite, C1 A/B against the warp-per-problem kernel
mkdir -p gpurun_out
TAG=${1:-c1}
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
cat > /tmp/cfg_bench.py <<'PY'
import sys, time, os
sys.path.insert(0, os.getcwd())
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
which = sys.argv[1]
pb = {"c1": lambda: synth.config1(batch=4096), "c2": lambda: synth.config2(batch=4096, horizon=100)}[which]()
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
    print("%s ms %.3f solves/s %.0f backward %.3f forward %.3f" % (which, dt * 1e3, pb.batch / dt, pr["riccati_backward"][0], pr["linesearch_rollout"][0]))
PY
{
for W in c1 c2; do
  echo col; python /tmp/cfg_bench.py $W
  echo "col groups=1"; ILQRB_GROUPS=1 python /tmp/cfg_bench.py $W
  echo "col groups=3"; ILQRB_GROUPS=3 python /tmp/cfg_bench.py $W
  echo warp; ILQRB_BWD=warp python /tmp/cfg_bench.py $W
done
} > gpurun_out/${TAG}_ab.txt 2>&1
cat gpurun_out/${TAG}_ab.txt
