#!/bin/bash
# Batch-solver Cholesky on the GPU box: batch-solver tests, TMA against the register-staged kernel, ncu of the factorisation.
mkdir -p gpurun_out
TAG=${1:-r2c}
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_n3.py -m gpu -q -x -k "batch or c4 or quad" > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
cat > /tmp/chol_bench.py <<'PY'
import sys, time
sys.path.insert(0, ".")
import numpy as np
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle
B = int(sys.argv[1]) if len(sys.argv) > 1 else 592
pb = synth.config4(batch=B, horizon=200)
u0 = synth.u0_time(pb).reshape(B, -1)
with Handle(pb) as h:
    h.solve_batch(u0, 1)
    h.set_profiling(True)
    t0 = time.perf_counter()
    out = h.solve_batch(u0, 3)
    dt = time.perf_counter() - t0
    print("B %d ms/iter %.2f" % (B, dt * 1e3 / 3), {k: round(v[0], 2) for k, v in h.profile().items()}, "cost", float(out.cost.mean()), float(np.abs(out.U).sum()))
PY
for e in "ILQRB_CHOL_TMA_WARPS=16" "ILQRB_CHOL_TMA_WARPS=8" "ILQRB_CHOL_NO_TMA=1"; do echo "env $e"; env $e timeout 600 python /tmp/chol_bench.py 592; env $e timeout 600 python /tmp/chol_bench.py 1024; done > gpurun_out/${TAG}_sweep.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_batch_chol -c 1 -o gpurun_out/${TAG}_chol -f python /tmp/chol_bench.py 148 > gpurun_out/${TAG}_ncu_f.log 2>&1
ncu -i gpurun_out/${TAG}_chol.ncu-rep --page raw --csv > gpurun_out/${TAG}_chol_raw.csv 2>/dev/null
tail -4 gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_sweep.txt
