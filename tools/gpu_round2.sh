#!/bin/bash
# Full single-GPU recipe of round 2: tests, smoke, default bench line (+ per-layer log), reference arm, training-step profile
# (ReLU bit mask A/B), launch list with DRAM bytes for profiles/r02_traffic.json.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --layers-log gpurun_out/layers.log > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/bench_n1.json; tail -5 gpurun_out/bench_n1.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/bench_ref.json
HPE_B200_BN_BITS=1 timeout 300 python tools/profile_train.py 256 > gpurun_out/profile_train_bits1.log 2>&1; head -4 gpurun_out/profile_train_bits1.log
HPE_B200_BN_BITS=0 timeout 300 python tools/profile_train.py 256 > gpurun_out/profile_train_bits0.log 2>&1; head -4 gpurun_out/profile_train_bits0.log
grep "bn_" gpurun_out/profile_train_bits1.log | cut -c1-90; grep "bn_" gpurun_out/profile_train_bits0.log | cut -c1-90
HPE_B200_CUDA_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'conv_gemm|yhalo|preprocess|decode' -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --quick --steps 2 --warmup 3 --no-cpu-baseline --sustain-s 0 > gpurun_out/ncu_list.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_list.log | cut -c1-300
# DINOv3Pose inference / training profiles and the ncu --set full capture of the chained kernel (scratch .ncu-rep, summary in profiles/)
timeout 600 python tools/bench_pose.py small 256 > gpurun_out/pose_profile.log 2>&1; head -3 gpurun_out/pose_profile.log | cut -c1-160
timeout 600 python tools/bench_posetrain.py > gpurun_out/posetrain_profile.log 2>&1; head -2 gpurun_out/posetrain_profile.log | cut -c1-160
timeout 300 python tools/time_preprocess.py > gpurun_out/preprocess_timing.log 2>&1; cat gpurun_out/preprocess_timing.log
bash tools/profile_chain.sh
