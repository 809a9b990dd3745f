#!/bin/bash
# One 8-GPU box: the 1/2/4/8 weak-scaling curve of the train step (the driver's own command line), configs[2] at 8 GPUs, the
# sharded inference of configs[3] and [4], the 2-GPU collected tests.  Every bench.py run is timed by the wall clock as well.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
run() {   # name, nproc, extra args
  local name=$1 np=$2; shift 2
  local t0=$(date +%s)
  if [ "$np" = 1 ]; then timeout 300 python bench.py --gpus 1 "$@" > gpurun_out/s8_${name}.log 2> gpurun_out/s8_${name}.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) bench.py --gpus $np "$@" > gpurun_out/s8_${name}.log 2> gpurun_out/s8_${name}.err; fi
  local rc=$?
  echo "== $name rc=$rc wall=$(( $(date +%s) - t0 ))s $(grep '^{' gpurun_out/s8_${name}.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('value %.1f %s, %.3f ms/step, e2e %.1f' % (d['value'], d['unit'], d['ms_per_step'], d['e2e']['value']))" 2>/dev/null)"
}
timeout 400 python -m pytest tests -q -m gpu -k "two_gpu or torchrun" > gpurun_out/s8_tests2.log 2>&1; echo "2-GPU tests rc=$? $(tail -1 gpurun_out/s8_tests2.log)"
for n in 1 2 4 8; do run train_${n}gpu $n --steps 20 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer; done
run perceptual_8gpu 8 --steps 5 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer --perceptual 0.1
run infer512_8gpu 8 --steps 5 --warmup 3 --no-cpu-baseline --no-eager-baseline --mode infer --input 512 --batch 32
run infer320_8gpu 8 --steps 10 --warmup 3 --no-cpu-baseline --no-eager-baseline --mode infer --input 320 --batch 32
run infer256_8gpu 8 --steps 10 --warmup 3 --no-cpu-baseline --no-eager-baseline --mode infer
nvidia-smi topo -m > gpurun_out/s8_topo.txt 2>&1
NCCL_DEBUG=INFO timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29877 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-infer 2>&1 | grep -E "NVLS|Connected|via|Using network|nChannels|Ring|Tree" | head -20 > gpurun_out/s8_nccl_info.txt
