#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer --no-graph"
for k in "conv_first_kernel" "wgrad_first_kernel" "gn_finalize" "pack_" "gn_stats_kernel" "colsum" "final_head_kernel" "gn_apply_kernel" "head_blend_kernel" "upsample2x_slide" "ssim_fwd_maps"; do
  MRISR_BWD_SIDE=0 timeout 600 ncu --set full --clock-control none -k regex:$k -s 3 -c 2 -o gpurun_out/sm_${k} -f $B > gpurun_out/sm_ncu_${k}.log 2>&1
  ncu -i gpurun_out/sm_${k}.ncu-rep --page details --csv 2>/dev/null > gpurun_out/sm_${k}.details.csv
  rm -f gpurun_out/sm_${k}.ncu-rep
done
ls gpurun_out/sm_*.csv | wc -l
