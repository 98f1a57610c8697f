#!/bin/bash
# A/B of an environment switch on one box: conv + model parity tests with the default build, then bench.py (headline + per-layer log)
# once per value.  usage: tools/gpu_ab.sh VAR val1 val2 ...   (outputs under gpurun_out/)
set -u
mkdir -p gpurun_out
VAR=$1; shift
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_model_gpu.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_quick.log
for val in "$@"; do
  export $VAR=$val
  timeout 600 python bench.py --steps 20 --warmup 5 --no-train --no-hires --no-cpu-baseline --sustain-s 0 --layers-log gpurun_out/layers_${VAR}_$val.log > gpurun_out/bench_${VAR}_$val.json 2> gpurun_out/bench_${VAR}_$val.err; echo "bench $VAR=$val rc=$?"; tail -3 gpurun_out/bench_${VAR}_$val.err
  python - <<PY
import json
j=json.load(open('gpurun_out/bench_${VAR}_$val.json'))
print('$VAR=$val', j['value'], j['ms_per_step'], j['roofline']['ms_forward'], 'e2e', j['e2e']['value'])
print({k:v['us_per_step'] for k,v in j['roofline']['per_kernel'].items()})
PY
done
