cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_conv_tc.py > gpurun_out/r12_tests.log 2>&1; echo rc=$? >> gpurun_out/r12_tests.log
tail -5 gpurun_out/r12_tests.log
MRISR_GNBWD_FUSED=0 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r12_bench_off.log 2>&1
cut -c1-200 gpurun_out/r12_bench_off.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r12_layers.txt > gpurun_out/r12_bench.log 2>&1
cut -c1-200 gpurun_out/r12_bench.log
grep gn_bwd gpurun_out/r12_layers.txt
for kb in 60 190; do
MRISR_GNBWD_SMEM_KB=$kb timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r12_layers_$kb.txt > gpurun_out/r12_bench_$kb.log 2>&1
cut -c1-200 gpurun_out/r12_bench_$kb.log
grep gn_bwd gpurun_out/r12_layers_$kb.txt
done
MRISR_GNBWD_CS=8 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r12_layers_cs8.txt > gpurun_out/r12_bench_cs8.log 2>&1
cut -c1-200 gpurun_out/r12_bench_cs8.log
grep gn_bwd gpurun_out/r12_layers_cs8.txt
