#!/bin/bash
# one ncu --set full capture of one kernel inside a short bench run:  scripts/ncu_one.sh <tag> <kernel regex> [skip] [count]
TAG=$1; KREGEX=$2; SKIP=${3:-2}; COUNT=${4:-1}
CMD="python bench.py --steps 1 --warmup 1 --passes 1 --subs 0 --torch-e2e 0 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$KREGEX" -s $SKIP -c $COUNT -o gpurun_out/${TAG} -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
