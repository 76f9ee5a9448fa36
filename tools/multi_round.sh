#!/bin/bash
# N-GPU visit: the in-process multi-GPU tests and the bench line at N ranks (torchrun, one rank per GPU).
N=${1:-2}
TAG=${2:-r2m}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity.py -m gpu -q -x -k "multi or woodbury" > gpurun_out/${TAG}_pytest_${N}gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_${N}gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err
echo "bench rc=$?"
tail -3 gpurun_out/${TAG}_pytest_${N}gpu.log; cut -c1-600 gpurun_out/${TAG}_bench_${N}gpu.json
