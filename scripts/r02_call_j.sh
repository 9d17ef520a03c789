#!/bin/bash
# matcher tests first (a pipeline deadlock must not hold the box), then the suite and the bench line
TAG=${1:-r02s}
timeout 300 python -m pytest tests/test_gpu_matcher_mma.py tests/test_gpu_matcher.py tests/test_gpu_comm.py -x -q -m gpu > gpurun_out/${TAG}_tests_matcher.log 2>&1; rc=$?; echo "matcher tests rc=$rc"
tail -5 gpurun_out/${TAG}_tests_matcher.log
if [ $rc -eq 0 ]; then bash scripts/r02_call_h.sh $TAG; fi
