cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r16_tests.log 2>&1; echo rc=$? >> gpurun_out/r16_tests.log
tail -5 gpurun_out/r16_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r16_layers.txt > gpurun_out/r16_bench.log 2>&1
cut -c1-200 gpurun_out/r16_bench.log
grep -E "l1_ssim|upsample2x_bwd" gpurun_out/r16_layers.txt
