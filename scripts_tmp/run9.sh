cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r33_tests.log 2>&1; echo rc=$? >> gpurun_out/r33_tests.log
tail -4 gpurun_out/r33_tests.log
MRISR_CONV_TC_MERGE3_BK32=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r33_layers_off.txt > gpurun_out/r33_bench_off.log 2>&1
cut -c1-200 gpurun_out/r33_bench_off.log
grep -E "conv_tc:k3:(32>|64>32)" gpurun_out/r33_layers_off.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r33_layers.txt > gpurun_out/r33_bench.log 2>&1
cut -c1-200 gpurun_out/r33_bench.log
grep -E "conv_tc:k3:(32>|64>32)" gpurun_out/r33_layers.txt
