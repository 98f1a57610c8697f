#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dwconv7_tiled -s 2 -c 1 -f -o gpurun_out/dw384 python tools/run_dwconv.py 384 16 12 256 5 > gpurun_out/ncu_dw.log 2>&1; echo rc=$?
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dwconv7_tiled -s 2 -c 1 -f -o gpurun_out/dw96 python tools/run_dwconv.py 96 64 48 256 5 > gpurun_out/ncu_dw2.log 2>&1; echo rc=$?
