cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r25_tests.log 2>&1; echo rc=$? >> gpurun_out/r25_tests.log
tail -4 gpurun_out/r25_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r25_layers.txt > gpurun_out/r25_bench.log 2>&1
cut -c1-1200 gpurun_out/r25_bench.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/r25_bench_infer.log 2>&1
cut -c1-900 gpurun_out/r25_bench_infer.log
