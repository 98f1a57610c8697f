timeout 400 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_n1.json
timeout 200 python tools/microbench.py forward 256 > gpurun_out/per_layer_v3.log 2>&1; head -1 gpurun_out/per_layer_v3.log
timeout 300 python bench.py --height 384 --width 288 --batch 64 --steps 30 --no-cpu-baseline > gpurun_out/bench_384x288_b64.json 2> gpurun_out/bench_384.err; echo "rc=$?"; cut -c1-200 gpurun_out/bench_384x288_b64.json
timeout 300 python bench.py --mode pose --steps 20 --no-cpu-baseline > gpurun_out/bench_pose.json 2> gpurun_out/bench_pose.err; echo "rc=$?"; cut -c1-200 gpurun_out/bench_pose.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
