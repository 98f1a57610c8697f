#!/bin/bash
# A/B of one environment switch on the training step: tools/gpu_train_ab.sh VAR "v1 v2" REPS
set -u
mkdir -p gpurun_out
VAR=$1; VALS=$2; REPS=${3:-2}
timeout 1200 python -m pytest tests/test_train_gpu.py -x -q -m gpu > gpurun_out/pytest_train.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_train.log
for rep in $(seq 1 $REPS); do for v in $VALS; do
  env $VAR=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-hires --no-pose-train --no-cpu-baseline --sustain-s 0 > gpurun_out/bench_train_${VAR}_${v}_$rep.json 2> gpurun_out/bench_train.err || tail -3 gpurun_out/bench_train.err
  python -c "
import json
j=json.load(open('gpurun_out/bench_train_${VAR}_${v}_$rep.json'))
t=j['train']; print('$VAR=$v rep $rep: train', t['value'], 'img/s', t['ms_per_step'], 'ms', t.get('clocks',{}).get('sm_mhz'))"
done; done
