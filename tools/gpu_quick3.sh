#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_posemodel_gpu.py tests/test_train_gpu.py tests/test_wgrad_gpu.py tests/test_dgrad_gpu.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_quick.log
timeout 600 python tools/profile_train.py 256 > gpurun_out/profile_train.log 2>&1; cat gpurun_out/profile_train.log | cut -c1-170
