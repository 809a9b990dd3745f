#!/bin/bash
# ncu --set full of every gn_bwd launch of one train step (all kinds), details page as CSV
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer --no-graph"
MRISR_BWD_SIDE=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:gn_bwd_fused_kernel -s 60 -c 20 -o gpurun_out/gnb -f $B > gpurun_out/gnb_ncu.log 2>&1
ncu -i gpurun_out/gnb.ncu-rep --page details --csv 2>/dev/null > gpurun_out/gnb.details.csv
ncu -i gpurun_out/gnb.ncu-rep --page raw --csv 2>/dev/null > gpurun_out/gnb.raw.csv
ls -la gpurun_out/gnb.*
rm -f gpurun_out/gnb.ncu-rep
