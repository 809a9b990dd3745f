cd $GRAFT_REPO_ROOT
( time timeout 900 python bench.py ) > gpurun_out/r15_bench_default.log 2>&1
cut -c1-1500 gpurun_out/r15_bench_default.log | tail -5
( time timeout 900 python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/r15_bench_ref.log 2>&1
tail -6 gpurun_out/r15_bench_ref.log | cut -c1-900
( time timeout 600 python bench.py --mode infer --steps 5 --warmup 3 --no-cpu-baseline ) > gpurun_out/r15_bench_infer.log 2>&1
tail -6 gpurun_out/r15_bench_infer.log | cut -c1-700
( time timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r15_smoke.log 2>&1
tail -8 gpurun_out/r15_smoke.log
nproc; lscpu | grep "Model name"
