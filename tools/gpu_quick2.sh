#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_posemodel_gpu.py tests/test_train_gpu.py tests/test_loss_gpu.py tests/test_model_gpu.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_quick.log
timeout 300 python tools/microbench.py loss > gpurun_out/loss_sweep.log 2>&1; tail -12 gpurun_out/loss_sweep.log
