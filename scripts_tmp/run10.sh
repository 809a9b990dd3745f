cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu -k "wgrad_rows" > gpurun_out/r17_tests.log 2>&1; echo rc=$? >> gpurun_out/r17_tests.log
tail -25 gpurun_out/r17_tests.log
