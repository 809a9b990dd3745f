cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r23_tests.log 2>&1; echo rc=$? >> gpurun_out/r23_tests.log
tail -3 gpurun_out/r23_tests.log
MRISR_RING_CTAS=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r23_layers_1.txt > gpurun_out/r23_bench_1.log 2>&1
cut -c1-200 gpurun_out/r23_bench_1.log
grep conv_ring gpurun_out/r23_layers_1.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r23_layers_2.txt > gpurun_out/r23_bench_2.log 2>&1
cut -c1-200 gpurun_out/r23_bench_2.log
grep conv_ring gpurun_out/r23_layers_2.txt
