#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py -q -m gpu > gpurun_out/pl_tests.log 2>&1; echo "pipeline tests rc=$? $(tail -1 gpurun_out/pl_tests.log)"
grep -E "^E  |FAILED|Error" gpurun_out/pl_tests.log | head -40
timeout 300 python tools/pipeline_bench.py > gpurun_out/pl_bench.log 2>&1; tail -20 gpurun_out/pl_bench.log
