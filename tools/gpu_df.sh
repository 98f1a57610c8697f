#!/bin/bash
# depth-first section sweep: chunk size x number of stages (bench headline only), after the model parity tests
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_model_gpu.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_quick.log
for cfg in "0 0" "32 1" "32 2" "64 2" "16 2" "64 3" "32 3"; do
  set -- $cfg
  export HPE_B200_DF_CHUNK=$1 HPE_B200_DF_STAGES=$2
  timeout 600 python bench.py --steps 20 --warmup 5 --no-train --no-hires --no-cpu-baseline --sustain-s 0 --layers-log gpurun_out/layers_df_$1_$2.log > gpurun_out/bench_df_$1_$2.json 2> gpurun_out/bench_df_$1_$2.err; echo "bench chunk=$1 stages=$2 rc=$?"; grep -v WARNING gpurun_out/bench_df_$1_$2.err | tail -3
  python - <<PY
import json
j=json.load(open('gpurun_out/bench_df_$1_$2.json'))
print('chunk=$1 stages=$2', j['value'], j['ms_per_step'], j['roofline']['ms_forward'], 'e2e', j['e2e']['value'], 'launches', j['gpu_launches_per_step'])
PY
done
