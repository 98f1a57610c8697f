#!/bin/bash
# Are layer1/layer2's 1x1 convs HBM-bound?  Timings with the inputs flushed from / resident in L2, then ncu --set full captures.
set -u
mkdir -p gpurun_out
for nf in 0 1; do for B in 256 32 16; do for e in 0 1; do
  RUN_LAYER_NOFLUSH=$nf HPE_B200_EPI2=$e timeout 120 python tools/run_layer.py 0 $B 64 48 64 256 1 1 1 8 2>&1 | tail -1
done; done; done
for nf in 0 1; do for B in 256 32 16; do
  RUN_LAYER_NOFLUSH=$nf timeout 120 python tools/run_layer.py 0 $B 64 48 256 64 1 1 0 8 2>&1 | tail -1
  RUN_LAYER_NOFLUSH=$nf timeout 120 python tools/run_layer.py 1 $B 64 48 64 64 1 1 0 8 2>&1 | tail -1
done; done
NCU="ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s 2 -c 1 -f"
HPE_B200_EPI2=0 timeout 300 $NCU -o gpurun_out/l1conv3 python tools/run_layer.py 0 256 64 48 64 256 1 1 1 5 > gpurun_out/ncu_l1a.log 2>&1; echo rc=$?
HPE_B200_EPI2=1 timeout 300 $NCU -o gpurun_out/l1conv3_epi2 python tools/run_layer.py 0 256 64 48 64 256 1 1 1 5 > gpurun_out/ncu_l1b.log 2>&1; echo rc=$?
timeout 300 $NCU -o gpurun_out/l1conv1 python tools/run_layer.py 0 256 64 48 256 64 1 1 0 5 > gpurun_out/ncu_l1c.log 2>&1; echo rc=$?
timeout 300 $NCU -o gpurun_out/l2conv2 python tools/run_layer.py 1 256 32 24 128 128 1 1 0 5 > gpurun_out/ncu_l2a.log 2>&1; echo rc=$?
HPE_B200_EPI2=0 timeout 300 $NCU -o gpurun_out/l2conv3 python tools/run_layer.py 0 256 32 24 128 512 1 1 1 5 > gpurun_out/ncu_l2b.log 2>&1; echo rc=$?
