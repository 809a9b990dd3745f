#!/bin/bash
# Round-2 evidence pass on one B200: GPU suite, smoke, the driver's bench line (train + infer sub-record + eager GPU baseline + CPU
# baseline), the other configurations, the reference arm, ncu launch list + traffic record, SASS instruction census.
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-r02}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$? $(tail -1 gpurun_out/${tag}_tests.log)"
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -4 gpurun_out/${tag}_smoke.log
timeout 900 python bench.py --layer-report gpurun_out/${tag}_layers.txt > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; grep '^{' gpurun_out/${tag}_bench.log | tail -1 | cut -c1-200
timeout 600 python bench.py --no-cpu-baseline --mode infer > gpurun_out/${tag}_bench_infer.log 2>&1; grep '^{' gpurun_out/${tag}_bench_infer.log | tail -1 | cut -c1-160
timeout 600 python bench.py --no-cpu-baseline --no-eager-baseline --mode infer --input 512 --batch 64 --steps 5 --layer-report gpurun_out/${tag}_layers_infer512.txt > gpurun_out/${tag}_bench_infer512.log 2>&1; grep '^{' gpurun_out/${tag}_bench_infer512.log | tail -1 | cut -c1-160
timeout 600 python bench.py --no-cpu-baseline --no-eager-baseline --mode infer --input 320 --batch 32 > gpurun_out/${tag}_bench_infer320.log 2>&1; grep '^{' gpurun_out/${tag}_bench_infer320.log | tail -1 | cut -c1-160
timeout 600 python bench.py --no-cpu-baseline --no-eager-baseline --no-infer --perceptual 0.1 --steps 5 --layer-report gpurun_out/${tag}_layers_perceptual.txt > gpurun_out/${tag}_bench_perceptual.log 2>&1; grep '^{' gpurun_out/${tag}_bench_perceptual.log | tail -1 | cut -c1-160
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_ref.log 2>&1; grep '^{' gpurun_out/${tag}_bench_ref.log | tail -1 | cut -c1-200
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer --no-graph"
MRISR_BWD_SIDE=0 timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size \
  --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_list.log 2>&1
python tools/ncu_launch_list.py gpurun_out/${tag}_launches.csv --json gpurun_out/${tag}_traffic.json > gpurun_out/${tag}_launches_step.md 2> gpurun_out/${tag}_launches_step.err; tail -1 gpurun_out/${tag}_launches_step.err; head -8 gpurun_out/${tag}_launches_step.md | cut -c1-150
for k in "ssim_grad_kernel" "gn_bwd_fused_kernel" "wgrad_rows_kernel" "conv_ring_kernel" "conv_tc_kernel"; do
  MRISR_BWD_SIDE=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 3 -o gpurun_out/${tag}_full_${k} -f $B > gpurun_out/${tag}_ncu_${k}.log 2>&1
  ncu -i gpurun_out/${tag}_full_${k}.ncu-rep --page details --csv 2>/dev/null > gpurun_out/${tag}_full_${k}.details.csv; ls -la gpurun_out/${tag}_full_${k}.details.csv | awk '{print $5, $9}'
  rm -f gpurun_out/${tag}_full_${k}.ncu-rep
done
