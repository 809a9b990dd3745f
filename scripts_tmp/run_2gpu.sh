cd $GRAFT_REPO_ROOT
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r17_bench_2gpu.log 2>&1; echo rc=$? >> gpurun_out/r17_bench_2gpu.log
grep -E "^\{|rc=" gpurun_out/r17_bench_2gpu.log | cut -c1-400
tail -5 gpurun_out/r17_bench_2gpu.log | cut -c1-300
