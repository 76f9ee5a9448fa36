for i in 1 2; do
python bench.py --steps 10 --warmup 3 --no-configs 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('new', d['value'], d['kernel_ms_per_solve']['riccati_backward'])"
(cd _ab && python bench.py --steps 10 --warmup 3 --no-configs 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('old', d['value'], d['kernel_ms_per_solve']['riccati_backward'])")
done
nvidia-smi -q -d POWER,CLOCK,PERFORMANCE | grep -i -A3 "power limit\|Throttle\|Reasons\|SM  " | head -60
