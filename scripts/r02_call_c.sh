#!/bin/bash
# FAST v2 (persistent, TMA-staged): parity first, then the bench line and an e2e host-chunk sweep
TAG=${1:-r02c}
timeout 600 python -m pytest tests/test_gpu_detector.py tests/test_gpu_sweep.py -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests.log
tail -8 gpurun_out/${TAG}_tests.log
if grep -q "rc=0" gpurun_out/${TAG}_tests.log; then
  for hc in 167 125 250; do
    python bench.py --steps 3 --warmup 3 --subs 0 --no-cpu --torch-e2e 0 --host-chunk $hc > gpurun_out/${TAG}_bench_hc$hc.json 2> gpurun_out/${TAG}_bench_hc$hc.err; echo "bench hc=$hc rc=$?"
  done
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests_all.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests_all.log
  tail -4 gpurun_out/${TAG}_tests_all.log
fi
