#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/na_alone.log 2>&1; echo "alone rc=$? $(tail -1 gpurun_out/na_alone.log)"
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/na_all.log 2>&1; echo "all rc=$? $(tail -1 gpurun_out/na_all.log)"; grep -E "FAILED|Error" gpurun_out/na_all.log | head
grep -E "no_alloc" gpurun_out/na_all.log | head -3
compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/na_memcheck.log 2>&1; echo "memcheck rc=$? $(grep -E 'ERROR SUMMARY' gpurun_out/na_memcheck.log | tail -1)"
