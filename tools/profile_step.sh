#!/bin/bash
# Round profile recipe (run on the GPU box): tests, bench, ncu launch list of one step, one --set full capture.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_n1.json
HPE_B200_CUDA_GRAPH=0 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:'conv_gemm|yhalo|maxpool|preprocess|decode' -c 240 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 120 python tools/exp_stem.py > gpurun_out/stem.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:yhalo -s 10 -c 1 -o gpurun_out/stem_pool_yhalo -f \
  python tools/exp_stem.py > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; cat gpurun_out/stem.log
