#!/bin/bash
# one ncu --set full capture of a kernel (regex $1) of the workload $2 (profile_run.py name), tag $3
mkdir -p gpurun_out
K=$1; WL=${2:-c2}; TAG=${3:-one}
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed
ILQRB_GROUPS=1 timeout 600 ncu --set full --metrics $M --clock-control none --import-source on -k regex:$K -s 1 -c 1 -o gpurun_out/${TAG} -f python tools/profile_run.py 3 $WL > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i gpurun_out/${TAG}.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
tail -3 gpurun_out/${TAG}_ncu.log
