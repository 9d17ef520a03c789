#!/bin/bash
# tcgen05 Hamming experiment: small case first (a pipeline deadlock must not hold the box), then the bench size
cd scripts/exp
timeout 25 ./mma_hamming 8 2000 > ../../gpurun_out/mma_small.log 2>&1; echo "rc=$?" >> ../../gpurun_out/mma_small.log
if grep -q "OK (bit-exact)" ../../gpurun_out/mma_small.log; then
  timeout 60 ./mma_hamming 999 2000 > ../../gpurun_out/mma_full.log 2>&1; echo "rc=$?" >> ../../gpurun_out/mma_full.log
fi
tail -12 ../../gpurun_out/mma_small.log; tail -22 ../../gpurun_out/mma_full.log 2>/dev/null
