#!/bin/bash
# ncu evidence for the timed build: launch list of one train step (-> profiles/rNN_launches_step.md, rNN_traffic.json) and
# --set full captures of the dominant kernels.  `gpurun --timeout 2400 -- 'bash tools/r02_ncu.sh TAG'`
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-r02}
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer --no-graph"
MRISR_BWD_SIDE=0 $B > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
MRISR_BWD_SIDE=0 timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size \
  --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_list.log 2>&1
wc -l gpurun_out/${tag}_launches.csv
python tools/ncu_launch_list.py gpurun_out/${tag}_launches.csv --json gpurun_out/${tag}_traffic.json > gpurun_out/${tag}_launches_step.md 2> gpurun_out/${tag}_launches_step.err; tail -2 gpurun_out/${tag}_launches_step.err; head -30 gpurun_out/${tag}_launches_step.md
for k in "ssim_" "gn_bwd_fused_kernel" "wgrad_rows_kernel" "conv_ring_kernel"; do
  n=$(echo $k | tr -d '_')
  MRISR_BWD_SIDE=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 6 -o gpurun_out/${tag}_full_${n} -f $B > gpurun_out/${tag}_ncu_${n}.log 2>&1
  ls -la gpurun_out/${tag}_full_${n}.ncu-rep 2>/dev/null | awk '{print $5, $9}'
done
