cd $GRAFT_REPO_ROOT
for m in 0 1 2; do
echo "=== base offset mode $m"
MRISR_TC_BASEOFF=$m timeout 300 python -m pytest tests/test_gpu_conv_tc.py -q -m gpu -k "forward_engines or dgrad_engines" 2>&1 | grep "passed\|failed\|AssertionError: (" | head -12
done
