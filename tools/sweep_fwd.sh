#!/bin/bash
# forward-pass variants of the single-integrator fast path on the bench workload: value (solves/s) and per-kernel ms per solve
mkdir -p gpurun_out
run() { env "$@" python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('%-28s value %8.0f e2e %8.0f  %s' % ('$*', d['value'], d['e2e']['value'], {k: round(v,3) for k,v in d['kernel_ms_per_solve'].items()}))"; }
{
run ILQRB_FWD_V1=1
run ILQRB_FWD_SPLIT=1
run ILQRB_FWD_VARIANT=0
run ILQRB_FWD_VARIANT=1
run ILQRB_FWD_VARIANT=2
run ILQRB_FWD_VARIANT=0 ILQRB_GROUPS=1
run ILQRB_FWD_VARIANT=0 ILQRB_GROUPS=2
run ILQRB_FWD_VARIANT=0 ILQRB_GROUPS=4
} > gpurun_out/sweep_fwd.txt 2>&1
cat gpurun_out/sweep_fwd.txt
