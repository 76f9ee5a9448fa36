#!/bin/bash
# e2e (host buffers) against the number of problem groups: streaming, blocking and no-gains forms of bench.py
mkdir -p gpurun_out
TAG=${1:-e2e}
for G in default 1 2 3 4 5 8; do
  if [ "$G" == "default" ]; then unset ILQRB_GROUPS; else export ILQRB_GROUPS=$G; fi
  python bench.py --steps 10 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('G=$G value %.0f e2e %.0f sync %.0f nogains %.0f' % (d['value'], d['e2e']['value'], d['e2e']['sync_call']['value'], d['e2e']['no_gains']['value']))"
done > gpurun_out/${TAG}_groups.txt 2>&1
cat gpurun_out/${TAG}_groups.txt
