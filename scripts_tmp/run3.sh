cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r10_tests.log 2>&1; echo rc=$? >> gpurun_out/r10_tests.log
tail -3 gpurun_out/r10_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r10_layers.txt > gpurun_out/r10_bench.log 2>&1
cut -c1-300 gpurun_out/r10_bench.log
grep "gn_bwd\|upsample2x_bwd\|gn_apply" gpurun_out/r10_layers.txt
