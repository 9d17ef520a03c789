#!/bin/bash
# fixed cost of a map-to-frame step by query count, then the launch list of the same script (per-kernel times)
timeout 300 python scripts/exp/map_fixed_cost.py > gpurun_out/r02_map_fixed.log 2>&1; echo "rc=$?"; cat gpurun_out/r02_map_fixed.log | tail -8
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"hamming|expand|merge" -c 400 --csv --log-file gpurun_out/r02_map_fixed_launches.csv python scripts/exp/map_fixed_cost.py > gpurun_out/r02_map_fixed_ncu.log 2>&1; echo "ncu rc=$?"
