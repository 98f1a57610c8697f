#!/bin/bash
# ncu --set full of the chained conv3+conv1 kernel (layer1.1.conv3 + residual -> layer1.2.conv1), B=256 @64x48
set -u
mkdir -p gpurun_out
timeout 120 python tools/run_layer.py 0 256 64 48 64 256 1 1 1 8 0 64 2>&1 | tail -1
timeout 120 python tools/run_layer.py 0 256 64 48 64 256 1 1 1 8 0 128 2>&1 | tail -1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s 2 -c 1 -f -o gpurun_out/chain64 python tools/run_layer.py 0 256 64 48 64 256 1 1 1 5 0 64 > gpurun_out/ncu_chain.log 2>&1; echo rc=$?
