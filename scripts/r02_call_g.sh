#!/bin/bash
# blur side-stream overlap experiment: same bench line with and without ORBX_BLUR_OVERLAP, then the detector / sequence parity tests with it on
TAG=${1:-r02j}
B="python bench.py --steps 3 --warmup 3 --subs 0 --no-cpu --torch-e2e 0 --stage-timers 0"
for v in 0 1 0 1; do
  ORBX_BLUR_OVERLAP=$v timeout 600 $B > gpurun_out/${TAG}_ov${v}.json 2> gpurun_out/${TAG}_ov${v}.err; echo "overlap=$v rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_ov${v}.json").read().strip().splitlines()[-1])
    print("  value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"])
except Exception as e: print("  parse failed", e)
PY
done
ORBX_BLUR_OVERLAP=1 timeout 900 python -m pytest tests/test_gpu_detector.py tests/test_gpu_native_sequence.py -m gpu -x -q > gpurun_out/${TAG}_tests_ov1.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/${TAG}_tests_ov1.log
