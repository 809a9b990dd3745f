cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu > gpurun_out/r11_tests.log 2>&1; echo rc=$? >> gpurun_out/r11_tests.log
tail -8 gpurun_out/r11_tests.log
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_conv_tc.py > gpurun_out/r11_tests2.log 2>&1; echo rc=$? >> gpurun_out/r11_tests2.log
tail -5 gpurun_out/r11_tests2.log
MRISR_WGRAD_TC3=0 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r11_bench_off.log 2>&1
cut -c1-200 gpurun_out/r11_bench_off.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r11_layers.txt > gpurun_out/r11_bench.log 2>&1
cut -c1-300 gpurun_out/r11_bench.log
head -50 gpurun_out/r11_layers.txt
