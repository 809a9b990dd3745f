#!/bin/bash
# Round-2 call 2: new parity tests on the promoted default build, smoke, bench with the new records, knob A/Bs under graph replay.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/c2_tests.log 2>&1; echo "tests rc=$? $(tail -1 gpurun_out/c2_tests.log)"
grep -E "FAILED|ERROR" gpurun_out/c2_tests.log | head -30
timeout 600 python -m pytest tests/test_gpu_layers.py tests/test_gpu_parity.py -q -m gpu -s -k "every_layer or full_size_train or trained_weights" > gpurun_out/c2_layers_verbose.log 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/c2_smoke.log 2>&1; tail -5 gpurun_out/c2_smoke.log
timeout 900 python bench.py --layer-report gpurun_out/c2_layers.txt > gpurun_out/c2_bench.log 2> gpurun_out/c2_bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/c2_bench.log | tail -1; tail -3 gpurun_out/c2_bench.err
bench() {
  local name=$1; shift
  env "$@" timeout 400 python bench.py --no-cpu-baseline --no-eager-baseline --no-infer --steps 10 --warmup 3 --layer-report gpurun_out/c2_${name}_layers.txt > gpurun_out/c2_${name}.log 2>&1
  local ms=$(tail -1 gpurun_out/c2_${name}.log | python -c "import sys,json; print('%.3f' % json.loads(sys.stdin.readline())['ms_per_step'])" 2>/dev/null)
  echo "== $name: ${ms:-FAILED} ms/step"
}
bench default MRISR_NONE=1
bench cs4 MRISR_GNBWD_CS=4
bench cs8 MRISR_GNBWD_CS=8
bench oneside MRISR_BWD_SIDE=0
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/c2_ref.log 2>&1; cut -c1-400 gpurun_out/c2_ref.log | tail -1
for f in gpurun_out/c2_*_layers.txt gpurun_out/c2_layers.txt; do echo "--- $f"; awk '{n=split($1,a,":"); fam[a[1]]+=$5} END{for(k in fam) printf "%s=%.3f ", k, fam[k]; print ""}' $f; done
