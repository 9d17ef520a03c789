#!/bin/bash
# First GPU call of the next round (one B200, ~3 min): what round 1 left unmeasured.
#   gpurun --timeout 400 -- 'bash scripts/next_round.sh r02'
# 1. the library-scheduled end-to-end path next to the Python-scheduled one (bench.py --native -> e2e_native)
# 2. launch list + ncu --set full of the pose kernels in their final configuration (8-hypothesis first wave,
#    interleaved bisection, two-SVD cheirality): profiles/r01_ncu_full_pose_v9.csv predates those changes
TAG=${1:-r02}
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests.log
python bench.py --native --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench_native.json 2> gpurun_out/${TAG}_bench_native.err
python bench.py --pose --frames 257 --chunk 257 --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1 && \
timeout 200 ncu --set full --clock-control none --import-source on -k regex:"pose_" -c 26 -o gpurun_out/${TAG}_pose_full -f \
  python bench.py --pose --frames 257 --chunk 257 --steps 1 --warmup 1 --no-cpu > gpurun_out/${TAG}_pose_ncu.log 2>&1
ncu -i gpurun_out/${TAG}_pose_full.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,smsp__inst_executed.sum > gpurun_out/${TAG}_pose_full_summary.csv 2>/dev/null
tail -3 gpurun_out/${TAG}_tests.log
