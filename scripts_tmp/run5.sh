cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu -k "vgg or perceptual" > gpurun_out/r12_tests.log 2>&1; echo rc=$? >> gpurun_out/r12_tests.log
tail -40 gpurun_out/r12_tests.log
