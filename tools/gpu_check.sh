#!/bin/bash
# Full GPU check used during development: `gpurun -- 'bash tools/gpu_check.sh TAG'`
# 1. the GPU test suite, 2. the train bench with the per-layer timing table, 3. the inference bench.
# Outputs land in gpurun_out/TAG_*.
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-check}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_tests.log 2>&1; echo rc=$? >> gpurun_out/${tag}_tests.log
tail -4 gpurun_out/${tag}_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/${tag}_layers.txt > gpurun_out/${tag}_bench.log 2>&1
cut -c1-400 gpurun_out/${tag}_bench.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/${tag}_bench_infer.log 2>&1
cut -c1-300 gpurun_out/${tag}_bench_infer.log
