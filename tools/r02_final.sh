#!/bin/bash
# Final pass of the round on one B200: GPU suite, ncu launch list of the shipped build (-> profiles/r02_traffic.json, read by bench.py
# when its source hash matches), then the driver's own command so that the recorded line carries roofline.traffic.
cd "${GRAFT_REPO_ROOT:-.}"
tag=${1:-r02f}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$? $(tail -1 gpurun_out/${tag}_tests.log)"
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer --no-graph"
MRISR_BWD_SIDE=0 timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size \
  --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_list.log 2>&1
python tools/ncu_launch_list.py gpurun_out/${tag}_launches.csv --json gpurun_out/${tag}_traffic.json > gpurun_out/${tag}_launches_step.md 2> gpurun_out/${tag}_launches_step.err; tail -1 gpurun_out/${tag}_launches_step.err
cp gpurun_out/${tag}_traffic.json profiles/r02_traffic.json
timeout 900 python bench.py --layer-report gpurun_out/${tag}_layers.txt > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
grep '^{' gpurun_out/${tag}_bench.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline'])"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_ref.log 2>&1; grep '^{' gpurun_out/${tag}_bench_ref.log | tail -1 | cut -c1-160
# base_filters = 64 (the default of the reference's inference command): side records
timeout 600 python bench.py --base-filters 64 --mode infer --no-cpu-baseline --no-eager-baseline > gpurun_out/${tag}_bench_f64_infer.log 2>&1; grep '^{' gpurun_out/${tag}_bench_f64_infer.log | tail -1 | cut -c1-170
timeout 600 python bench.py --base-filters 64 --no-cpu-baseline --no-eager-baseline --no-infer --steps 5 --layer-report gpurun_out/${tag}_layers_f64.txt > gpurun_out/${tag}_bench_f64_train.log 2>&1; grep '^{' gpurun_out/${tag}_bench_f64_train.log | tail -1 | cut -c1-170
