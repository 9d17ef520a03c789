#!/bin/bash
# round 2, GPU call: full GPU test-suite on the new tree, then the bench line (matcher auto = tensor cores) and the POPC one
TAG=${1:-r02b}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests.log
tail -15 gpurun_out/${TAG}_tests.log
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --steps 3 --warmup 3 --matcher popc --subs 0 --no-cpu > gpurun_out/${TAG}_bench_popc.json 2> gpurun_out/${TAG}_bench_popc.err; echo "bench popc rc=$?"
tail -3 gpurun_out/${TAG}_bench.err
