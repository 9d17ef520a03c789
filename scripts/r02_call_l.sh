#!/bin/bash
# FAST without the inverted planes: detector parity first, then the whole suite and a bench line
TAG=${1:-r02x}
timeout 600 python -m pytest tests/test_gpu_detector.py tests/test_gpu_sweep.py -x -q -m gpu > gpurun_out/${TAG}_tests_det.log 2>&1; rc=$?; echo "detector tests rc=$rc"; tail -3 gpurun_out/${TAG}_tests_det.log
if [ $rc -eq 0 ]; then
  timeout 900 python bench.py --steps 3 --warmup 3 --subs 0 --no-cpu --torch-e2e 0 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
  python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_bench.json"))
print("value %.0f e2e %.0f" % (j["value"], j["e2e"]["value"]), {k:round(v["ms_per_pass"],3) for k,v in j["stages"].items()})
PY
fi
