#!/bin/bash
# Round-end style check: `gpurun -- 'bash tools/gpu_full.sh TAG'`
# GPU tests, smoke, default bench (with cpu_baseline), inference benches (256^2, 512^2, 320^2), the perceptual configuration,
# the reference arm; with NCU=1 also one `ncu --set full` capture of the dominant elementwise and conv kernels.
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-full}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_tests.log 2>&1; echo rc=$? >> gpurun_out/${tag}_tests.log
tail -4 gpurun_out/${tag}_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -3 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py --layer-report gpurun_out/${tag}_layers.txt > gpurun_out/${tag}_bench.log 2>&1
cut -c1-400 gpurun_out/${tag}_bench.log | tail -1
timeout 600 python bench.py --no-cpu-baseline --mode infer > gpurun_out/${tag}_bench_infer.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_infer.log | tail -1
timeout 600 python bench.py --no-cpu-baseline --mode infer --input 512 --batch 64 --steps 5 --layer-report gpurun_out/${tag}_layers_infer512.txt > gpurun_out/${tag}_bench_infer512.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_infer512.log | tail -1
timeout 600 python bench.py --no-cpu-baseline --mode infer --input 320 --batch 32 > gpurun_out/${tag}_bench_infer320.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_infer320.log | tail -1
timeout 600 python bench.py --no-cpu-baseline --perceptual 0.1 --steps 5 --layer-report gpurun_out/${tag}_layers_perceptual.txt > gpurun_out/${tag}_bench_perceptual.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_perceptual.log | tail -1
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_ref.log | tail -1
if [ -n "$NCU" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gn_bwd_fused_kernel -s 36 -c 6 -o gpurun_out/${tag}_gn_bwd -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${tag}_ncu_gn.log 2>&1
tail -1 gpurun_out/${tag}_ncu_gn.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size --clock-control none -s 592 -c 450 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${tag}_ncu_list.log 2>&1
wc -l gpurun_out/${tag}_launches.csv
fi
