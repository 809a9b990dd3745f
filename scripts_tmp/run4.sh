cd $GRAFT_REPO_ROOT
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r11_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gn_bwd_ -s 120 -c 8 -o gpurun_out/r11_gnbwd python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r11_ncu.log 2>&1
tail -3 gpurun_out/r11_ncu.log
