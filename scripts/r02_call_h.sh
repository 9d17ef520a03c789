#!/bin/bash
# wave-aware split choice of the tensor-core matcher + the configs[4] sub-measurement: suite, then the default bench line
TAG=${1:-r02k}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests_all.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/${TAG}_tests_all.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_bench.json"))
print("value %.0f e2e %.0f" % (j["value"], j["e2e"]["value"]))
for k in ("hd","seqs","map","latency"):
    print(k, {a:b for a,b in j.get(k,{}).items() if a not in ("workload","timer","api")})
PY
