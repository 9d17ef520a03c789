#!/bin/bash
# round 2, GPU call 1: state of the tree at round start + the experiments that need no new code
TAG=r02a
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests.log
python bench.py --native --steps 5 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench_native.json 2> gpurun_out/${TAG}_bench_native.err
for c in 64 128 250; do
python bench.py --chunk $c --steps 5 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench_chunk$c.json 2> gpurun_out/${TAG}_bench_chunk$c.err
done
ncu --query-metrics 2>/dev/null | grep -i -E "pipe_(xu|alu|int|uniform|fma)" > gpurun_out/${TAG}_ncu_pipe_metrics.txt
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
tail -3 gpurun_out/${TAG}_tests.log
