#!/bin/bash
# y-halo kernel: timing of the stem (plain / pooled) and layer1 conv2, then one --set full capture of the pooled stem
set -u
mkdir -p gpurun_out
for i in 1 2 3; do timeout 120 python tools/exp_stem.py 2>&1 | tail -3; done | tee gpurun_out/stem.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:yhalo -s 10 -c 1 -o gpurun_out/stem_pool_yhalo -f \
  python tools/exp_stem.py > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
