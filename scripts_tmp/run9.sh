cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r20_tests.log 2>&1; echo rc=$? >> gpurun_out/r20_tests.log
tail -4 gpurun_out/r20_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r20_layers.txt > gpurun_out/r20_bench.log 2>&1
cut -c1-200 gpurun_out/r20_bench.log
grep -E "head_blend|blend_dalpha|wgrad" gpurun_out/r20_layers.txt
