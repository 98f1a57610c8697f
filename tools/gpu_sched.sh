#!/bin/bash
# A/B of the L2 schedule (HPE_B200_L2_SCHED mode : HPE_B200_L2_KEEP_MB) on one box; each combination is benched `REPS` times.
# usage: tools/gpu_sched.sh "0:104 3:104 7:104 3:60" [REPS]
set -u
mkdir -p gpurun_out
REPS=${2:-2}
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_model_gpu.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_quick.log
for rep in $(seq 1 $REPS); do
for combo in $1; do
  export HPE_B200_L2_SCHED=${combo%%:*} HPE_B200_L2_KEEP_MB=${combo##*:}
  timeout 600 python bench.py --steps 20 --warmup 5 --no-train --no-hires --no-pose-train --no-cpu-baseline --sustain-s 0 > gpurun_out/bench_sched_${combo/:/_}_$rep.json 2> gpurun_out/bench_sched.err || tail -3 gpurun_out/bench_sched.err
  python - <<PY
import json
j=json.load(open('gpurun_out/bench_sched_${combo/:/_}_$rep.json'))
print('sched $combo rep $rep: fwd', j['roofline']['ms_forward'], 'step', j['ms_per_step'], 'e2e', j['e2e']['value'])
PY
done
done
# 192-wide tiles against wave quantisation (layer4 conv1 / conv2): parity of the affected layers, then the bench
if [ -n "${WAVE_AB:-}" ]; then
  unset HPE_B200_L2_SCHED HPE_B200_L2_KEEP_MB
  HPE_B200_WAVE192=1 timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_model_gpu.py -x -q 2>&1 | tail -3
  for rep in 1 2 3; do for v in 0 1; do
    HPE_B200_WAVE192=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-train --no-hires --no-pose-train --no-cpu-baseline --sustain-s 0 --layers-log gpurun_out/layers_wave$v.log > gpurun_out/bench_wave${v}_$rep.json 2> gpurun_out/bench_sched.err || tail -3 gpurun_out/bench_sched.err
    python -c "
import json
j=json.load(open('gpurun_out/bench_wave${v}_$rep.json'))
print('wave192=$v rep $rep: fwd', j['roofline']['ms_forward'], {k:v['us_per_step'] for k,v in j['roofline']['per_kernel'].items() if 'layer4' in k})"
  done; done
fi
