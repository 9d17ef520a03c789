#!/bin/bash
# one ncu --set full capture of every kernel of one 1000-frame device-resident pass (the second pass of a short run)
TAG=${1:-r02_ncu_full_v3}
CMD="python bench.py --steps 1 --warmup 1 --passes 1 --subs 0 --torch-e2e 0 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on \
  -k regex:"hamming_mma|expand_desc|merge_splits|cross_mask|compact_kernel|fast_nms|blur7|orient|harris|select|brief|resize" -s 19 -c 19 \
  -o gpurun_out/${TAG} -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
