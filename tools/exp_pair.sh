timeout 400 python -m pytest tests/test_wgrad_gpu.py tests/test_train_gpu.py tests/test_dgrad_gpu.py tests/test_conv_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --mode train --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_train_n1.json 2> gpurun_out/bench_train_n1.err; cut -c1-600 gpurun_out/bench_train_n1.json
