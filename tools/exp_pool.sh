#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -x -q -k "stem or maxpool or conv3x3" 2>&1 | tail -3
timeout 120 python tools/exp_stem.py 2>&1 | tail -3
HPE_B200_HALO_DIRECT=0 timeout 120 python tools/exp_stem.py 2>&1 | tail -3
HPE_B200_HALO_PROF=1 timeout 120 python tools/exp_stem.py 2>&1 | grep -v "^$" | awk 'NR%8==1 || /us/'
