#!/bin/bash
# quick GPU check of a kernel change: conv + model parity tests, then bench --quick with a per-layer log (A/B via env)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_model_gpu.py tests/test_dgrad_gpu.py tests/test_posemodel_gpu.py tests/test_train_gpu.py tests/test_loss_gpu.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_quick.log
for tag in direct tma; do
  if [ $tag = tma ]; then export HPE_B200_EPI_TMA=1; else unset HPE_B200_EPI_TMA; fi
  timeout 600 python bench.py --steps 20 --warmup 5 --no-train --no-hires --no-cpu-baseline --sustain-s 0 --layers-log gpurun_out/layers_$tag.log > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench $tag rc=$?"
  python - <<PY
import json
j=json.load(open('gpurun_out/bench_$tag.json'))
print('$tag', j['value'], j['ms_per_step'], j['roofline']['ms_forward'], 'e2e', j['e2e']['value'], j['e2e']['fp32_images']['value'])
print({k:v['us_per_step'] for k,v in j['roofline']['per_kernel'].items()})
PY
done
unset HPE_B200_EPI_TMA
timeout 300 python tools/microbench.py loss > gpurun_out/loss_sweep.log 2>&1; tail -12 gpurun_out/loss_sweep.log
