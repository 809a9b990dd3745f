cd $GRAFT_REPO_ROOT
for n in 8 4; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r17_bench_${n}gpu.log 2>&1; echo rc=$? >> gpurun_out/r17_bench_${n}gpu.log
grep -E "^\{|rc=" gpurun_out/r17_bench_${n}gpu.log | cut -c1-260
done
