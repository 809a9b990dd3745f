"""Data-parallel CUDA-graph check, run under torchrun on >= 2 GPUs:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_graph_dp.py
Three training steps with the eager data-parallel step and with GraphedTrainStep(buckets=...) from the same initial state must
give the same parameters on every rank (same kernels, same NCCL reductions, only launched differently).
CHECK_BASE_FILTERS=16 / 64 runs the same check at another width (another bucket plan)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from mri_superresolution_b200.models.unet_model import UNetSuperRes
from mri_superresolution_b200.optim import FusedAdam
from mri_superresolution_b200.parallel import GradientBuckets
from mri_superresolution_b200.utils.graph import GraphedTrainStep
from mri_superresolution_b200.utils.losses import CombinedLoss
from mri_superresolution_b200.utils.synthetic import make_pairs


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    say = lambda m: print(f"[check_graph_dp rank {rank}] {m}", flush=True)      # noqa: E731
    say("process group up")
    batches = [tuple(t.to(dev) for t in make_pairs(4, 64, 64, seed=100 + i, rank=rank)) for i in range(3)]
    results = []
    for use_graph in (False, True):
        torch.manual_seed(7)
        model = UNetSuperRes(base_filters=int(os.environ.get("CHECK_BASE_FILTERS", "32"))).to(dev)
        crit = CombinedLoss(ssim_weight=0.4)
        opt = FusedAdam(model, lr=1e-3, weight_decay=1e-5, capturable=use_graph)
        buckets = GradientBuckets(model)
        losses = []
        if use_graph:
            stepper = GraphedTrainStep(model, crit, opt, batches[0][0], batches[0][1], buckets=buckets)
            say("graph recorded")
        for x, t in batches:
            if use_graph:
                losses.append(float(stepper(x, t).item()))
            else:
                opt.zero_grad(set_to_none=True)
                loss = crit(model(x), t)
                loss.backward()
                opt.step(grad_scale=buckets.allreduce())
                losses.append(float(loss.item()))
        results.append((losses, model.flat_parameters().clone()))
        say(f"{'graph' if use_graph else 'eager'} leg done: {losses}")
        if use_graph:
            stepper.close()
        buckets.close()
    (l0, p0), (l1, p1) = results
    err = float((p0 - p1).abs().max() / p0.abs().max())
    # replicas must agree with each other, too
    ref = p1.clone()
    dist.broadcast(ref, 0)
    spread = float((p1 - ref).abs().max())
    print(f"rank {rank}: eager losses {l0} graph losses {l1} param rel err {err:.2e} replica spread {spread:.2e}", flush=True)
    ok = err < 1e-5 and spread == 0.0 and all(abs(a - b) <= 1e-6 * abs(a) for a, b in zip(l0, l1))
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    okay = int(flag.item()) == 1
    if rank == 0:
        print("GRAPH_DP_OK" if okay else "GRAPH_DP_MISMATCH", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    import threading
    wd = threading.Timer(20.0, lambda: os._exit(0 if okay else 1))
    wd.daemon = True
    wd.start()
    dist.destroy_process_group()
    wd.cancel()
    sys.exit(0 if okay else 1)


if __name__ == "__main__":
    main()
