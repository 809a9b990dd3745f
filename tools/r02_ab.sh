#!/bin/bash
# A/B of one or more environment settings against the default build on one box: tools/r02_ab.sh NAME "ENV=VAL ..." [NAME2 "..."] ...
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
bench() {
  local name=$1; shift
  env "$@" timeout 400 python bench.py --no-cpu-baseline --no-eager-baseline --no-infer --steps 10 --warmup 3 --layer-report gpurun_out/ab_${name}_layers.txt > gpurun_out/ab_${name}.log 2>&1
  local ms=$(grep '^{' gpurun_out/ab_${name}.log | tail -1 | python -c "import sys,json; print('%.3f' % json.loads(sys.stdin.readline())['ms_per_step'])" 2>/dev/null)
  echo "== $name: ${ms:-FAILED} ms/step"; [ -z "$ms" ] && tail -5 gpurun_out/ab_${name}.log
  awk '{n=split($1,a,":"); fam[a[1]]+=$5} END{for(k in fam) printf "%s=%.3f ", k, fam[k]; print ""}' gpurun_out/ab_${name}_layers.txt 2>/dev/null
}
if [ -n "$TESTS" ]; then timeout 1200 python -m pytest tests -x -q -m gpu $TESTS > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$? $(tail -1 gpurun_out/ab_tests.log)"; grep -E "^E  |FAILED" gpurun_out/ab_tests.log | head -20; fi
bench default MRISR_NONE=1
while [ $# -ge 2 ]; do bench "$1" $2; shift 2; done
bench default2 MRISR_NONE=1
