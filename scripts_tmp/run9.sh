cd $GRAFT_REPO_ROOT
nvidia-smi -L
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 ) > gpurun_out/r16_bench_2gpu.log 2>&1
tail -4 gpurun_out/r16_bench_2gpu.log | cut -c1-900
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 2 --warmup 1 --impl reference ) > gpurun_out/r16_bench_2gpu_ref.log 2>&1
tail -4 gpurun_out/r16_bench_2gpu_ref.log | cut -c1-400
