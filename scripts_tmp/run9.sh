cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r36_tests.log 2>&1; echo rc=$? >> gpurun_out/r36_tests.log
tail -4 gpurun_out/r36_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r36_layers.txt > gpurun_out/r36_bench.log 2>&1
cut -c1-200 gpurun_out/r36_bench.log
grep -E "gn_finalize" gpurun_out/r36_layers.txt
