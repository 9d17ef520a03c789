#!/bin/bash
# Round profile capture: bench line, ncu launch list of the same command, one --set full capture per kernel.
# usage: scripts/gpu_profile.sh <tag>   (outputs under gpurun_out/)
TAG=${1:-r01_v4}
set -x
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu.log 2>&1
ncu --set full --clock-control none --import-source on \
  -k regex:"hamming|fast_nms|blur7|orient|harris|select|brief|resize" -s 18 -c 16 -o gpurun_out/${TAG}_full -f \
  python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/${TAG}_ncu_full.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,smsp__inst_executed.sum,l1tex__t_sectors.sum,lts__t_sectors.sum > gpurun_out/${TAG}_full_summary.csv 2>/dev/null
ls -la gpurun_out | tail -8
