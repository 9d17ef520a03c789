#!/bin/bash
# CTA pairs with multicast B tiles: matcher tests first (a deadlock must not hold the box), then A/B through ORBX_MMA_NO_PAIRS
TAG=${1:-r02bb}
timeout 300 python -m pytest tests/test_gpu_matcher_mma.py tests/test_gpu_matcher.py tests/test_gpu_comm.py -x -q -m gpu > gpurun_out/${TAG}_tests_matcher.log 2>&1; rc=$?; echo "matcher tests rc=$rc"
tail -4 gpurun_out/${TAG}_tests_matcher.log
if [ $rc -eq 0 ]; then
  for v in 1 0; do
    if [ $v -eq 1 ]; then export ORBX_MMA_NO_PAIRS=1; else unset ORBX_MMA_NO_PAIRS; fi
    timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --torch-e2e 0 --hd-passes 4 --latency-iters 5 > gpurun_out/${TAG}_np$v.json 2> gpurun_out/${TAG}_np$v.err
    python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_np$v.json")); print("no_pairs=$v value %.0f e2e %.0f hamming %.3f ms map %.3f ms (%.2f T pairs/s) hd %.0f" % (j["value"], j["e2e"]["value"], j["stages"]["hamming"]["ms_per_pass"], j["map"]["kernel_ms"], j["map"]["pairs_per_s"]/1e12, j["hd"]["value"]))
PY
  done
fi
