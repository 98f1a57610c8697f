#!/bin/bash
# N-GPU box: the default bench line under torchrun (inference + e2e + hires + train + pose_train), nothing else
set -u
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n$N rc=$?"
python - <<PY
import json
j=json.load(open('gpurun_out/bench_n$N.json'))
print('value', j['value'], j['ms_per_step'], 'e2e', j['e2e']['value'], j['e2e']['ms_per_step'], 'fp32', j['e2e']['fp32_images']['value'])
for k in ('hires','train','pose_train'):
    print(k, j[k]['value'], j[k]['ms_per_step'])
PY
tail -3 gpurun_out/bench_n$N.err | cut -c1-300
