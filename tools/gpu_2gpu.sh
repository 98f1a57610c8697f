#!/bin/bash
# 2-GPU box: distributed tests (DP training step + clean teardown, 2-GPU bench line), then the full bench under torchrun
set -u
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_dist_gpu.py tests/test_posemodel_gpu.py tests/test_loss_gpu.py tests/test_decode_gpu.py -x -q > gpurun_out/pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_2gpu.log
timeout 300 python tools/microbench.py loss > gpurun_out/loss_sweep.log 2>&1; tail -12 gpurun_out/loss_sweep.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; cut -c1-600 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
j=json.load(open('gpurun_out/bench_n2.json'))
for k in ('value','ms_per_step','e2e','hires','train'):
    print(k, json.dumps(j.get(k))[:700])
PY
