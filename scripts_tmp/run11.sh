cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r18_tests.log 2>&1; echo rc=$? >> gpurun_out/r18_tests.log
tail -3 gpurun_out/r18_tests.log
for mb in 0 16 32 48 80; do
  MRISR_GNBWD_CHUNK_MB=$mb timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r18_bench_$mb.log 2>&1
  python - <<PY
import json
d=json.loads(open('gpurun_out/r18_bench_$mb.log').read().split('\n')[0])
k=[x for x in d['kernels'] if x['kernel']=='gn_bwd'][0]
print('chunk_mb', $mb, 'step ms', round(d['ms_per_step'],3), 'gn_bwd ms', k['ms_per_step'], 'launches', d['gpu_launches'])
PY
done
