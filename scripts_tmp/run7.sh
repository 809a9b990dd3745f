cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r14_tests.log 2>&1; echo rc=$? >> gpurun_out/r14_tests.log
tail -5 gpurun_out/r14_tests.log
