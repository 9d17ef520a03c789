#!/bin/bash
# compile-time parameter sweep on the GPU box: scripts/exp/param_sweep.sh  (results on stdout)
cd /root/repo
run() { python bench.py --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['value']), {k:round(v['ms_per_step'],3) for k,v in d['stages'].items() if k!='blur+brief'})"; }
try() { # file, sed-expr, label
  cp $1 /tmp/_orig.cu; sed -i "$2" $1; make -C slam-rs_b200/csrc >/dev/null 2>&1 && run "$3" || echo "$3 build failed"; cp /tmp/_orig.cu $1; }
run base
try slam-rs_b200/csrc/select.cu 's/constexpr int kHarrisBlocksPerLevel = 2;/constexpr int kHarrisBlocksPerLevel = 1;/' harris1
try slam-rs_b200/csrc/select.cu 's/constexpr int kOrientPerWarp = 16;/constexpr int kOrientPerWarp = 32;/' orient32
try slam-rs_b200/csrc/select.cu 's/constexpr int kHarrisScan = 128;/constexpr int kHarrisScan = 256;/' harrisscan256
try slam-rs_b200/csrc/match.cu 's/#pragma unroll 1\n    for (; j + 4 <= rows; j += 4) {/XX/' noop
make -C slam-rs_b200/csrc >/dev/null 2>&1
