#!/bin/bash
# usage: scripts/ncu_kernel.sh <kernel-regex> <tag> [skip]   -> gpurun_out/<tag>.ncu-rep + <tag>.txt (one launch, --set full)
K="$1"; TAG="$2"; SKIP="${3:-2}"
ncu --set full --clock-control none --import-source on -k regex:"$K" -s "$SKIP" -c 1 -o gpurun_out/$TAG -f \
  python bench.py --frames 250 --chunk 250 --steps 1 --warmup 1 --no-cpu > gpurun_out/$TAG.log 2>&1
ncu -i gpurun_out/$TAG.ncu-rep --page details > gpurun_out/$TAG.txt 2>&1
grep -E "Duration|Executed Ipc Active|Issue Slots Busy|Registers Per|Achieved Occupancy|Theoretical Occupancy|DRAM Throughput|L1/TEX Hit|Memory Throughput|No Eligible|Warp Cycles Per Issued|Compute \(SM\) Throughput" gpurun_out/$TAG.txt
