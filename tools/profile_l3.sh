#!/bin/bash
# ncu --set full of layer3's 1x1 convs (conv1: 1024->256, conv3: 256->1024 + residual), B=256 @16x12
set -u
mkdir -p gpurun_out
timeout 120 python tools/run_layer.py 0 256 16 12 1024 256 1 1 0 5 2>&1 | tail -1
timeout 120 python tools/run_layer.py 0 256 16 12 256 1024 1 1 1 5 2>&1 | tail -1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s 2 -c 1 -o gpurun_out/l3conv1 -f python tools/run_layer.py 0 256 16 12 1024 256 1 1 0 5 > gpurun_out/ncu_l3a.log 2>&1; echo rc=$?
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s 2 -c 1 -o gpurun_out/l3conv3 -f python tools/run_layer.py 0 256 16 12 256 1024 1 1 1 5 > gpurun_out/ncu_l3b.log 2>&1; echo rc=$?
