set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu > gpurun_out/r8_conv_tests.log 2>&1; echo rc=$? >> gpurun_out/r8_conv_tests.log
tail -5 gpurun_out/r8_conv_tests.log
for a in "64 512 512 16 16 fwd" "64 256 256 32 32 fwd" "64 256 256 64 32 fwd" "64 512 512 32 16 fwd" "64 256 256 32 64 fwd" "64 512 512 16 16 wgrad" "64 256 256 32 32 wgrad" "64 512 512 32 16 wgrad" "64 256 256 64 32 wgrad"; do
  PYTHONPATH=. timeout 120 python tests/prof_conv.py 2 $a 4 >> gpurun_out/r8_prof.log 2>&1
done
cat gpurun_out/r8_prof.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r8_parity.log 2>&1; echo rc=$? >> gpurun_out/r8_parity.log
tail -5 gpurun_out/r8_parity.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r8_layers.txt > gpurun_out/r8_bench.log 2>&1
cut -c1-400 gpurun_out/r8_bench.log
