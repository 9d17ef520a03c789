#!/bin/bash
# dress rehearsal of the round-end driver sequence on one B200: GPU tests, smoke(), both bench arms with the driver's flags
if [ "$1" != "bench" ]; then
timeout 1200 python -m pytest tests/ -x -q -m gpu > gpurun_out/reh_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/reh_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/reh_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/reh_smoke.log
fi
S=$SECONDS; timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/reh_ref.json 2> gpurun_out/reh_ref.err; echo "ref rc=$? wall $((SECONDS-S)) s"
S=$SECONDS; timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/reh_bench.json 2> gpurun_out/reh_bench.err; echo "bench rc=$? wall $((SECONDS-S)) s"
python - <<PY
import json
j=json.load(open("gpurun_out/reh_bench.json")); r=json.load(open("gpurun_out/reh_ref.json"))
print("ours value %.0f e2e %.0f ms_per_step %.2f steps %d launches %s clocks %s" % (j["value"], j["e2e"]["value"], j["ms_per_step"], j["steps"], j["gpu_launches"], j["clocks"]))
print("ref value %.2f %s steps %d; same config: %s" % (r["value"], r["unit"], r["steps"], j["config"] == r["config"]))
PY
