#!/bin/bash
# Train-step throughput against the per-GPU batch size (does a smaller batch, whose tensors fit the L2, run faster per slice?)
cd "${GRAFT_REPO_ROOT:-.}"
for b in 64 32 16 8 4; do
  python bench.py --batch $b --no-cpu-baseline --no-eager-baseline --no-infer --steps 20 2>/dev/null | grep '^{' | tail -1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('batch $b: %.0f slices/s, %.3f ms/step' % (d['value'], d['ms_per_step']))"
done
