#!/bin/bash
# streaming e2e of bench.py: chained handles (ilqrb_chain_after_compute) against unchained and against a start-up stagger, per group count
mkdir -p gpurun_out
run() { python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1 value %.0f e2e %.0f sync %.0f nogains %.0f' % (d['value'], d['e2e']['value'], d['e2e']['sync_call']['value'], d['e2e']['no_gains']['value']))"; }
{
run "default (chain once)"
for G in 1 2 3 5; do ILQRB_GROUPS=$G run "G=$G chain once"; done
ILQRB_BENCH_CHAIN=never run "default unchained"
ILQRB_BENCH_CHAIN=always ILQRB_GROUPS=2 run "G=2 chained always"
ILQRB_BENCH_CHAIN=never ILQRB_GROUPS=2 ILQRB_BENCH_STAGGER_MS=4 run "G=2 unchained stagger=4ms"
} > gpurun_out/e2e_chain.txt 2>&1
cat gpurun_out/e2e_chain.txt
