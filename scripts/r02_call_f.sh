#!/bin/bash
# comm with per-rank launch threads (virtual ranks on one GPU) + the whole suite, then the XU-pipe counter of the POPC kernel
TAG=${1:-r02i}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests_all.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests_all.log
tail -4 gpurun_out/${TAG}_tests_all.log
CMD="python bench.py --steps 1 --warmup 1 --passes 1 --subs 0 --torch-e2e 0 --no-cpu --matcher popc"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_alu.sum,smsp__inst_executed.sum,sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.sum.per_cycle_active \
  --clock-control none -k regex:"hamming_knn|popc_bench" -c 6 --csv --log-file gpurun_out/${TAG}_popc_xu.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
