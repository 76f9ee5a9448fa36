#!/bin/bash
# Column-per-lane backward kernel (fused1_col.cuh) on the GPU box: parity tests, A/B timing against the warp-per-problem kernel.
mkdir -p gpurun_out
TAG=${1:-r2c}
[ "$2" == "notest" ] || timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_n3.py tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
cat > /tmp/c2_bench.py <<'PY'
import sys, time
sys.path.insert(0, ".")
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
pb = synth.config2(batch=B, horizon=100)
u0 = None
with Handle(pb) as h:
    for rep in range(4):
        h.upload_u0(u0)
        h.synchronize()
        t0 = time.perf_counter()
        h.run_recursive(20, True, False)
        h.synchronize()
        dt = time.perf_counter() - t0
    h.set_profiling(True)
    h.upload_u0(u0)
    h.run_recursive(20, True, False)
    h.synchronize()
    pr = h.profile()
    print("B %d ms %.3f solves/s %.0f backward %.3f forward %.3f" % (B, dt * 1e3, pb.batch / dt, pr["riccati_backward"][0], pr["linesearch_rollout"][0]))
PY
for B in 4096 8192; do
  echo "c2"; timeout 300 python /tmp/c2_bench.py $B
done > gpurun_out/${TAG}_ab.txt 2>&1
cat gpurun_out/${TAG}_ab.txt
