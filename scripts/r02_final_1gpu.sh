#!/bin/bash
# final 1-GPU capture of the round: suite, both bench arms, then the ncu launch list of one short pass
TAG=${1:-r02_v3}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests_all.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests_all.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
SHORT="python bench.py --steps 1 --warmup 1 --passes 1 --subs 0 --torch-e2e 0 --no-cpu"
$SHORT > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $SHORT > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_bench.json"))
print("value %.0f e2e %.0f launches %s" % (j["value"], j["e2e"]["value"], j.get("gpu_launches")))
r=json.load(open("gpurun_out/${TAG}_bench_ref.json")); print("ref", r["value"], r["unit"])
PY
