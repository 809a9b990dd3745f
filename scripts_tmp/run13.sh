cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu > gpurun_out/r20_tests.log 2>&1; echo rc=$? >> gpurun_out/r20_tests.log
tail -15 gpurun_out/r20_tests.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r20_tests2.log 2>&1; echo rc=$? >> gpurun_out/r20_tests2.log
tail -4 gpurun_out/r20_tests2.log
for m in 0 1; do
MRISR_CONV_TC_MERGE3=$m timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r20_layers_$m.txt > gpurun_out/r20_bench_$m.log 2>&1
cut -c1-200 gpurun_out/r20_bench_$m.log
grep "conv_tc:k3" gpurun_out/r20_layers_$m.txt | head -8
done
