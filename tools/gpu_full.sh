#!/bin/bash
# Round-end style check: `gpurun -- 'bash tools/gpu_full.sh TAG'`
# GPU tests, smoke, default bench (with cpu_baseline), inference bench, reference arm, then one `ncu --set full`
# capture of the dominant elementwise kernel (gn_bwd_fused) and of the dominant conv kernel.
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-full}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_tests.log 2>&1; echo rc=$? >> gpurun_out/${tag}_tests.log
tail -4 gpurun_out/${tag}_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -3 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py --layer-report gpurun_out/${tag}_layers.txt > gpurun_out/${tag}_bench.log 2>&1
cut -c1-600 gpurun_out/${tag}_bench.log
timeout 600 python bench.py --no-cpu-baseline --mode infer > gpurun_out/${tag}_bench_infer.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_infer.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_ref.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gn_bwd_fused_kernel -s 36 -c 6 -o gpurun_out/${tag}_gn_bwd -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_gn.log 2>&1
tail -1 gpurun_out/${tag}_ncu_gn.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_ring_kernel -s 36 -c 6 -o gpurun_out/${tag}_conv_ring -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_ring.log 2>&1
tail -1 gpurun_out/${tag}_ncu_ring.log | cut -c1-200
ls -la gpurun_out
