#!/bin/bash
# One GPU call that times the unswept block-size knobs and the env-selected staged kernels against the default:
#   gpurun --timeout 1500 -- 'bash tools/sweep_knobs.sh'
# Each arm = the default train bench (batch 64, 256^2) with one setting changed; prints the step time and the affected
# per-family lines of the layer report.  Kernel variants that need another library are handled by tools/ab_variant.sh.
cd "${GRAFT_REPO_ROOT:-.}"
run() {   # name, grep pattern, env assignments...
  local name=$1 pat=$2; shift 2
  env "$@" timeout 400 python bench.py --no-cpu-baseline --steps 10 --warmup 3 --layer-report gpurun_out/sweep_${name}_layers.txt \
      > gpurun_out/sweep_${name}.log 2>&1
  local ms=$(tail -1 gpurun_out/sweep_${name}.log | python -c "import sys,json; print('%.3f' % json.loads(sys.stdin.readline())['ms_per_step'])" 2>/dev/null)
  echo "== $name: ${ms:-FAILED} ms/step"
  grep -h "$pat" gpurun_out/sweep_${name}_layers.txt 2>/dev/null | awk '{printf "   %-44s %s ms\n", $1, $5}' | head -8
}
run default "gn_apply\|conv_first\|l1_ssim\|head_blend" MRISR_NONE=1
run default_again "gn_apply\|conv_first\|l1_ssim\|head_blend" MRISR_NONE=1
for u in 32 64 128; do run gnapply_upt$u "gn_apply:pool0" MRISR_GNAPPLY_UPT=$u; done
for u in 8 16; do run gnapply_pool_upt$u "gn_apply:pool1" MRISR_GNAPPLY_POOL_UPT=$u; done
for u in 64 256; do run convfirst_upt$u "conv_first" MRISR_CONVFIRST_UPT=$u; done
for u in 8 32; do run headblend_upt$u "head_blend" MRISR_HEADBLEND_UPT=$u; done
# GroupNorm backward: do the L2-sized alternatives look different now that the step replays as one CUDA graph (no launch gaps)?
# (eager measurements: profiles/r01_gn_bwd_chunking.md, r01_gn_bwd_fused.md - chunking and clusters > 2 both lost)
for mb in 48 96; do run gnbwd_chunk${mb} "gn_bwd:kind0" MRISR_GNBWD_FUSED=0 MRISR_GNBWD_CHUNK_MB=$mb; done
for cs in 4 8; do run gnbwd_cs${cs} "gn_bwd:kind0" MRISR_GNBWD_CS=$cs; done
# the env-selected staged paths, each only after its parity test passes
if MRISR_TEST_STAGED=1 timeout 300 python -m pytest tests/test_gpu_knobs.py -q -k staged_ssim > gpurun_out/sweep_ssim_split_test.log 2>&1; then
  run ssim_split "l1_ssim" MRISR_SSIM_SPLIT=1
else
  echo "== ssim_split: parity test failed (gpurun_out/sweep_ssim_split_test.log)"; tail -5 gpurun_out/sweep_ssim_split_test.log
fi
if MRISR_TEST_STAGED=1 timeout 300 python -m pytest tests/test_gpu_knobs.py -q -k staged_side > gpurun_out/sweep_bwd_side_test.log 2>&1; then
  run bwd_side "wgrad\|conv_tc:k3:256" MRISR_BWD_SIDE=1
  MRISR_BWD_SIDE=1 timeout 300 python -m pytest tests/test_gpu_parity.py -q -k "graph" > gpurun_out/sweep_bwd_side_graph_test.log 2>&1
  echo "   graph tests under MRISR_BWD_SIDE=1: $(tail -1 gpurun_out/sweep_bwd_side_graph_test.log)"
else
  echo "== bwd_side: parity test failed (gpurun_out/sweep_bwd_side_test.log)"; tail -5 gpurun_out/sweep_bwd_side_test.log
fi
# library variants: bit-level comparison with the default on the probe shape (the full parity suite + timing: tools/ab_variant.sh)
MRISR_TEST_STAGED=1 timeout 600 python -m pytest tests/test_gpu_knobs.py -q -k staged_library > gpurun_out/sweep_libs_test.log 2>&1
echo "== staged libraries vs default: $(tail -1 gpurun_out/sweep_libs_test.log)"
