#!/bin/bash
# A/B of a staged build against the measured default: `gpurun -- 'bash tools/ab_variant.sh incr'` after `bash tools/build_variants.sh`
# (run here; the .so files travel with the snapshot).  Variants: incr (gn_bwd constant-step addressing), ups (sliding bilinear x2
# forward + scatter-form adjoint), dalpha (d(alpha) sums inside gn_bwd instead of blend_dalpha_kernel), all (everything).
# Runs the GPU parity tests under the variant first (a variant that fails them is not timed), then the default bench twice
# around the variant's (same box, back to back), with per-layer reports.
cd "${GRAFT_REPO_ROOT:-.}"
v=${1:?variant name}
lib=mri_superresolution_b200/lib/libmrisr_b200_${v}.so
[ -f "$lib" ] || { echo "missing $lib"; exit 2; }
MRISR_B200_VARIANT=$v timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/ab_${v}_tests.log 2>&1; rc=$?
tail -3 gpurun_out/ab_${v}_tests.log
[ $rc -eq 0 ] || { echo "variant $v fails the GPU tests (rc=$rc): not timed"; exit 1; }
for arm in default $v default2; do
  var=$v; [ "$arm" = "$v" ] || var=""
  MRISR_B200_VARIANT=$var timeout 600 python bench.py --no-cpu-baseline --layer-report gpurun_out/ab_${v}_${arm}_layers.txt \
      > gpurun_out/ab_${v}_${arm}_bench.log 2>&1
  echo "$arm: $(cut -c1-160 gpurun_out/ab_${v}_${arm}_bench.log | tail -1)"
done
if [ "$v" = ups ] || [ "$v" = all ]; then      # the forward kernel with the next source row prefetched one step ahead
  MRISR_B200_VARIANT=$v MRISR_UPSLIDE_PF=1 timeout 600 python bench.py --no-cpu-baseline --layer-report gpurun_out/ab_${v}_pf_layers.txt \
      > gpurun_out/ab_${v}_pf_bench.log 2>&1
  echo "$v + MRISR_UPSLIDE_PF=1: $(cut -c1-160 gpurun_out/ab_${v}_pf_bench.log | tail -1)"
  grep -h "upsample2x " gpurun_out/ab_${v}_pf_layers.txt | head -2
fi
grep -h "gn_bwd\|upsample\|blend" gpurun_out/ab_${v}_default_layers.txt | head -24
grep -h "gn_bwd\|upsample\|blend" gpurun_out/ab_${v}_${v}_layers.txt | head -24
