#!/bin/bash
# One GPU-box visit: parity suite, bench line, ncu launch list + one full capture of the C2 kernels.  Logs under gpurun_out/.
mkdir -p gpurun_out
TAG=${1:-r2}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q -s --durations=15 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
if [ "$2" != "noncu" ]; then
  M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_tensor_subpipe_dmma.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active
  ILQRB_GROUPS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv python tools/profile_run.py 3 c2 > gpurun_out/${TAG}_ncu_l.log 2>&1
  ILQRB_GROUPS=1 timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:k_\(backward_n1\|forward_ring_n1\) -s 3 -c 3 -o gpurun_out/${TAG}_c2 -f python tools/profile_run.py 3 c2 > gpurun_out/${TAG}_ncu_f.log 2>&1
  ncu -i gpurun_out/${TAG}_c2.ncu-rep --page raw --csv > gpurun_out/${TAG}_c2_raw.csv 2>/dev/null
  ILQRB_GROUPS=1 timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:k_\(backward_time\|forward_time\|kp_derivs_time\) -s 3 -c 3 -o gpurun_out/${TAG}_time -f python tools/profile_run.py 3 time > gpurun_out/${TAG}_ncu_t.log 2>&1
  ncu -i gpurun_out/${TAG}_time.ncu-rep --page raw --csv > gpurun_out/${TAG}_time_raw.csv 2>/dev/null
  python tools/ncu_kernels_json.py gpurun_out/${TAG}_c2_raw.csv 4096 "gpurun_out/${TAG}_c2_raw.csv: ncu --set full --clock-control none, ILQRB_GROUPS=1 python tools/profile_run.py 3 c2 (B=4096, one launch = whole batch)" gpurun_out/${TAG}_ncu_kernels.json > /dev/null 2>&1
fi
tail -5 gpurun_out/${TAG}_pytest.log
