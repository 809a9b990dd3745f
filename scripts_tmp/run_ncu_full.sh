cd $GRAFT_REPO_ROOT
timeout 1200 ncu --set full --import-source on --clock-control none --kernel-name regex:"conv_ring_kernel|wgrad_rows_kernel" --launch-skip 54 --launch-count 18 -o gpurun_out/r21_ring_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r21_ncu.log 2>&1
tail -2 gpurun_out/r21_ncu.log | cut -c1-200
ls -la gpurun_out/*.ncu-rep
