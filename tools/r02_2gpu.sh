#!/bin/bash
# 2-GPU teardown check: the collected 2-GPU test, the graph-vs-eager data-parallel check, one bench.py --gpus 2 run with its wall time.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "two_gpu" > gpurun_out/g2_test.log 2>&1; echo "2gpu test rc=$? $(tail -1 gpurun_out/g2_test.log)"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_graph_dp.py > gpurun_out/g2_graph_dp.log 2>&1; echo "check_graph_dp rc=$? $(grep -E 'GRAPH_DP' gpurun_out/g2_graph_dp.log | tail -1)"
t0=$(date +%s); timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/g2_bench.log 2> gpurun_out/g2_bench.err; echo "bench2 rc=$? wall=$(( $(date +%s)-t0 ))s"; grep '^{' gpurun_out/g2_bench.log | tail -1 | cut -c1-300
