#!/bin/bash
# N-GPU check (N = number of visible GPUs): orbx_comm tests with real ranks, then the bench line under torchrun
TAG=${1:-r02m}; N=${2:-2}
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_comm.py -x -q > gpurun_out/${TAG}_comm_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_comm_tests.log
tail -5 gpurun_out/${TAG}_comm_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/${TAG}_bench_${N}gpu.err
python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_bench_${N}gpu.json"))
print("value %.0f e2e %.0f n_gpus %d" % (j["value"], j["e2e"]["value"], j["n_gpus"]))
for k in ("hd","seqs","map","latency"):
    print(k, {a:b for a,b in j.get(k,{}).items() if a not in ("workload","timer","api")})
PY
