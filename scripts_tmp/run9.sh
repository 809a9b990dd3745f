cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_knobs.py -q -rP -m gpu > gpurun_out/r28_tests.log 2>&1; echo rc=$? >> gpurun_out/r28_tests.log
grep -E "passed|failed|^[a-z_0-9]+ \{|Assertion"  gpurun_out/r28_tests.log
