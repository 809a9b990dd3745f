cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r32_tests.log 2>&1; echo rc=$? >> gpurun_out/r32_tests.log
tail -4 gpurun_out/r32_tests.log
MRISR_UPBWD_TILED=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r32_layers_off.txt > gpurun_out/r32_bench_off.log 2>&1
cut -c1-200 gpurun_out/r32_bench_off.log
grep -E "upsample2x_bwd" gpurun_out/r32_layers_off.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r32_layers.txt > gpurun_out/r32_bench.log 2>&1
cut -c1-200 gpurun_out/r32_bench.log
grep -E "upsample2x_bwd" gpurun_out/r32_layers.txt
