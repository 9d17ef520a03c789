#!/bin/bash
# pipelined run_host passes with even chunks: parity tests, then the e2e figure for several host_chunk sizes
TAG=${1:-r02m}
timeout 600 python -m pytest tests/test_gpu_native_sequence.py -x -q -m gpu 2>&1 | tail -5
for hc in 84 125 167 250 500; do
  timeout 600 python bench.py --steps 3 --warmup 3 --subs 0 --no-cpu --torch-e2e 0 --host-chunk $hc > gpurun_out/${TAG}_hc${hc}.json 2> gpurun_out/${TAG}_hc${hc}.err; echo "host_chunk=$hc rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_hc${hc}.json").read().strip().splitlines()[-1])
    print("  value %.0f e2e %.0f ms/pass %.3f" % (d["value"], d["e2e"]["value"], d["e2e"]["ms_per_pass"]))
except Exception as e: print("  parse failed", e)
PY
done
