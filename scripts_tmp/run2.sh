cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r9_tests.log 2>&1; echo rc=$? >> gpurun_out/r9_tests.log
tail -5 gpurun_out/r9_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r9_layers.txt > gpurun_out/r9_bench.log 2>&1
cut -c1-300 gpurun_out/r9_bench.log
head -40 gpurun_out/r9_layers.txt
