#!/bin/bash
# Round-2 profile capture (one B200): the bench line + reference arm, the ncu launch list of one pass, one ncu --set full
# capture of every kernel of the 1000-frame device-resident pass.  usage: scripts/gpu_profile_r02.sh <tag>
TAG=${1:-r02_v2}
SHORT="python bench.py --steps 1 --warmup 1 --passes 1 --subs 0 --torch-e2e 0 --no-cpu"
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err; echo "ref rc=$?"
$SHORT > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $SHORT > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
$SHORT > gpurun_out/${TAG}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on \
  -k regex:"hamming|expand_desc|merge_splits|cross_mask|compact|fast_nms|blur7|orient|harris|select|brief|resize" -c 19 \
  -o gpurun_out/${TAG}_full -f $SHORT > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full rc=$?"
ls -la gpurun_out | grep ${TAG}
