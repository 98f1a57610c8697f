#!/bin/bash
# interleaved A/B of one environment switch with long timed regions: tools/gpu_ab2.sh VAR "v1 v2" REPS STEPS [pytest files]
set -u
mkdir -p gpurun_out
VAR=$1; VALS=$2; REPS=${3:-3}; STEPS=${4:-100}; shift 4
if [ $# -gt 0 ]; then timeout 1200 python -m pytest "$@" -x -q -m gpu > gpurun_out/pytest_ab2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_ab2.log; fi
for rep in $(seq 1 $REPS); do for v in $VALS; do
  env $VAR=$v timeout 600 python bench.py --steps $STEPS --warmup 10 --no-train --no-hires --no-pose-train --no-cpu-baseline --sustain-s 0 > gpurun_out/bench_${VAR}_${v}_$rep.json 2> gpurun_out/bench_ab2.err || tail -3 gpurun_out/bench_ab2.err
  python -c "
import json
j=json.load(open('gpurun_out/bench_${VAR}_${v}_$rep.json'))
print('$VAR=$v rep $rep: fwd', j['roofline']['ms_forward'], 'step', j['ms_per_step'], 'e2e', j['e2e']['value'], j['clocks']['sm_mhz'])"
done; done
