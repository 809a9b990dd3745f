cd $GRAFT_REPO_ROOT
run() {
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layer-report gpurun_out/r26_layers_$name.txt > gpurun_out/r26_bench_$name.log 2>&1
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/r26_bench_$name.log') if l.startswith('{')][-1])
    g=[k for k in d['kernels'] if k['kernel']=='gn_bwd'][0]
    print('$name', 'step', round(d['ms_per_step'],3), 'gn_bwd', g['ms_per_step'])
except Exception as e:
    print('$name', 'FAILED', e)
PY
}
MRISR_GNBWD_L2HINT=1 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r26_tests_h.log 2>&1; echo rc=$? >> gpurun_out/r26_tests_h.log
tail -3 gpurun_out/r26_tests_h.log
run cs2_h MRISR_GNBWD_CS=2 MRISR_GNBWD_L2HINT=1
run cs8_h MRISR_GNBWD_CS=8 MRISR_GNBWD_L2HINT=1
