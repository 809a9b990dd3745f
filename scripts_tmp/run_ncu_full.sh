cd $GRAFT_REPO_ROOT
timeout 1200 ncu --set full --import-source on --clock-control none --kernel-name regex:"upsample2x_bwd_tiled_kernel|l1_ssim_kernel|gn_bwd_fused_kernel" --launch-skip 69 --launch-count 23 -o gpurun_out/r30_elem_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r30_ncu.log 2>&1
tail -2 gpurun_out/r30_ncu.log | cut -c1-200
ls -la gpurun_out/r30*.ncu-rep
