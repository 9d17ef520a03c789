#!/bin/bash
# FAST variant check: parity (detector + sweep), one bench line, then the whole GPU suite
TAG=${1:-r02f}
timeout 600 python -m pytest tests/test_gpu_detector.py tests/test_gpu_sweep.py -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests.log
tail -8 gpurun_out/${TAG}_tests.log
if grep -q "rc=0" gpurun_out/${TAG}_tests.log; then
  python bench.py --steps 3 --warmup 3 --subs 0 --no-cpu --torch-e2e 0 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
  python -c "
import json; j=json.load(open('gpurun_out/${TAG}_bench.json'))
print('value %.0f e2e %.0f' % (j['value'], j['e2e']['value'])); print({k: round(v['ms_per_pass'], 3) for k, v in j['stages'].items()})"
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests_all.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests_all.log
  tail -4 gpurun_out/${TAG}_tests_all.log
fi
