#!/bin/bash
# ncu launch list of one train step: `gpurun -- 'bash tools/gpu_ncu.sh TAG'`, then `python tools/ncu_launch_list.py gpurun_out/TAG_launches.csv`
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-ncu}
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size --clock-control none -s 592 -c 450 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu.log 2>&1
tail -2 gpurun_out/${tag}_ncu.log | cut -c1-200
wc -l gpurun_out/${tag}_launches.csv
