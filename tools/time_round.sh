#!/bin/bash
# Time-system fast path on the GPU box: its tests, the ring-shape sweep of k_forward_time, ncu launch list + full capture.
mkdir -p gpurun_out
TAG=${1:-r2t}
timeout 1200 python -m pytest tests/test_gpu_hooks.py tests/test_pylqr.py tests/test_gpu_parity.py -m gpu -q -x -k "eval_point or debug_api or tutorial or time or indefinite or pivoted" > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
cat > /tmp/time_bench.py <<'PY'
import sys, time
sys.path.insert(0, ".")
import torch
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
pb = synth.config4(batch=4096, horizon=100)
u0 = synth.u0_time(pb)
with Handle(pb) as h:
    for rep in range(3):
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
    print("ms %.3f solves/s %.0f" % (dt * 1e3, pb.batch / dt), h.profile())
PY
for v in 0; do echo "variant $v"; ILQRB_TIME_FWD=$v timeout 300 python /tmp/time_bench.py; done > gpurun_out/${TAG}_sweep.txt 2>&1
for g in 1 2 3; do echo "groups $g"; ILQRB_GROUPS=$g timeout 300 python /tmp/time_bench.py; done >> gpurun_out/${TAG}_sweep.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv python tools/profile_run.py 3 time > gpurun_out/${TAG}_ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_\(backward_time\|forward_time\) -s 2 -c 2 -o gpurun_out/${TAG}_time -f python tools/profile_run.py 3 time > gpurun_out/${TAG}_ncu_f.log 2>&1
ncu -i gpurun_out/${TAG}_time.ncu-rep --page raw --csv > gpurun_out/${TAG}_time_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_time.ncu-rep --page source --csv > gpurun_out/${TAG}_time_source.csv 2>/dev/null
tail -5 gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_sweep.txt
