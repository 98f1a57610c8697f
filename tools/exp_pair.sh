timeout 240 python -m pytest tests/test_conv_gpu.py tests/test_dgrad_gpu.py -m gpu -x -q 2>&1 | tail -4
for pair in 0 1; do
HPE_B200_PAIR=$pair timeout 120 python tools/microbench.py forward 256 > gpurun_out/fwd_elect_pair$pair.log 2>&1; echo "pair=$pair $(head -1 gpurun_out/fwd_elect_pair$pair.log)"
done
