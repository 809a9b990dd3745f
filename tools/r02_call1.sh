#!/bin/bash
# Round-2 first GPU call: settle the staged variants of round 1 (parity first, then A/B timing), reduced knob sweep.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/c1_tests_default.log 2>&1; echo "default tests rc=$? $(tail -1 gpurun_out/c1_tests_default.log)"
bench() {  # name, env...
  local name=$1; shift
  env "$@" timeout 400 python bench.py --no-cpu-baseline --no-eager-baseline --no-infer --steps 10 --warmup 3 --layer-report gpurun_out/c1_${name}_layers.txt \
      > gpurun_out/c1_${name}.log 2>&1
  local ms=$(tail -1 gpurun_out/c1_${name}.log | python -c "import sys,json; print('%.3f' % json.loads(sys.stdin.readline())['ms_per_step'])" 2>/dev/null)
  echo "== $name: ${ms:-FAILED} ms/step"
}
bench default MRISR_NONE=1
for v in incr ups dalpha all; do
  MRISR_B200_VARIANT=$v timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/c1_tests_${v}.log 2>&1; rc=$?
  echo "variant $v tests rc=$rc $(tail -1 gpurun_out/c1_tests_${v}.log)"
  [ $rc -eq 0 ] && bench $v MRISR_B200_VARIANT=$v
done
bench default2 MRISR_NONE=1
bench ups_pf MRISR_B200_VARIANT=ups MRISR_UPSLIDE_PF=1
for u in 32 64; do bench gnapply_upt$u MRISR_GNAPPLY_UPT=$u; done
bench gnapply_pool_upt8 MRISR_GNAPPLY_POOL_UPT=8
for u in 64 256; do bench convfirst_upt$u MRISR_CONVFIRST_UPT=$u; done
for u in 8; do bench headblend_upt$u MRISR_HEADBLEND_UPT=$u; done
if MRISR_TEST_STAGED=1 timeout 300 python -m pytest tests/test_gpu_knobs.py -q -k staged_ssim > gpurun_out/c1_ssim_split_test.log 2>&1; then
  bench ssim_split MRISR_SSIM_SPLIT=1
else echo "ssim_split parity FAILED"; tail -5 gpurun_out/c1_ssim_split_test.log; fi
if MRISR_TEST_STAGED=1 timeout 300 python -m pytest tests/test_gpu_knobs.py -q -k staged_side > gpurun_out/c1_bwd_side_test.log 2>&1; then
  bench bwd_side MRISR_BWD_SIDE=1
  MRISR_BWD_SIDE=1 timeout 300 python -m pytest tests/test_gpu_parity.py -q -k "graph" > gpurun_out/c1_bwd_side_graph_test.log 2>&1
  echo "   graph tests under MRISR_BWD_SIDE=1: $(tail -1 gpurun_out/c1_bwd_side_graph_test.log)"
else echo "bwd_side parity FAILED"; tail -5 gpurun_out/c1_bwd_side_test.log; fi
bench default3 MRISR_NONE=1
for f in gpurun_out/c1_*_layers.txt; do echo "--- $f"; awk '{n=split($1,a,":"); fam[a[1]]+=$5} END{for(k in fam) printf "%s=%.3f ", k, fam[k]; print ""}' $f; done
