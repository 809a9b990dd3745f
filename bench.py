#!/usr/bin/env python
"""Benchmark of the U-Net super-resolution training step (BASELINE.json configs[1]) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode train|infer]

One "step" = one pass of the hot path over one batch of synthetic slices:
  train: zero_grad -> UNetSuperRes forward -> CombinedLoss(L1 + 0.4 SSIM) -> backward -> (bucketed NCCL
         gradient all-reduce when N > 1) -> fused Adam   (the loop body of scripts/train.py:296-318)
  infer: no_grad forward + clamp(0,1)                     (scripts/infer.py:268-276)
Per-GPU batch is fixed (64 slices of 256x256 -> 512x512), so N > 1 is weak scaling; ranks are launched by
torch.distributed.run (one process per GPU, NCCL).  Rank 0 prints ONE JSON line.

  value : slices/s with inputs resident in HBM (CUDA events, max over ranks)
  e2e   : slices/s through the public module API with inputs in pinned HOST memory, the host->device copies and
          the loss read-back (loss.item(), as scripts/train.py:321 does) inside the timed region
  roofline : the dominant kernel family of the step, timed live with CUDA events on the launching stream
             (mrisr_prof_*), algorithmic FLOPs (or bytes) / time against MEASURED_PEAKS.json
  cpu_baseline : the oracle (plain PyTorch fp32 restatement of the reference, oracle/) timed on the host cores on a
             bounded sample of the same workload.  `--impl reference` prints that arm alone.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BASE_FILTERS = 32
H = W = 256                     # network input (LR) size; target / output is 512 x 512
PER_GPU_BATCH = 64
SSIM_WEIGHT = 0.4
LR, WD = 1e-4, 1e-5             # scripts/train.py:186 defaults (lr 1e-4, weight_decay 1e-5)
SEED = 1234
METRIC = "U-Net train slices/sec at 256^2 bf16"
INFER_METRIC = "U-Net infer slices/sec at 256^2 bf16"


# ----------------------------------------------------------------------------------------- FLOP / byte model
def unet_conv_table(f: int):
    """(ks, cin, cout, level) of the 21 convolutions; level l -> (H >> l), -1 -> 2H."""
    return [(3, 1, f, 0), (3, f, f, 0), (3, f, 2 * f, 1), (3, 2 * f, 2 * f, 1), (3, 2 * f, 4 * f, 2),
            (3, 4 * f, 4 * f, 2), (3, 4 * f, 8 * f, 3), (3, 8 * f, 8 * f, 3),
            (1, 8 * f, 4 * f, 2), (3, 8 * f, 4 * f, 2), (3, 4 * f, 4 * f, 2),
            (1, 4 * f, 2 * f, 1), (3, 4 * f, 2 * f, 1), (3, 2 * f, 2 * f, 1),
            (1, 2 * f, f, 0), (3, 2 * f, f, 0), (3, f, f, 0),
            (3, f, f // 2, -1), (3, f, 2 * f, 0), (3, f // 2, f // 2, -1), (1, f // 2, 1, -1)]


def unet_flops_per_slice(f: int, h: int, w: int, train: bool) -> float:
    """2*MAC of the reference's 21 convs (1x1 up-convs counted at the reference's high resolution)."""
    fwd = 0.0
    first = 0.0
    for i, (ks, ci, co, lvl) in enumerate(unet_conv_table(f)):
        hh, ww = (2 * h, 2 * w) if lvl < 0 else (h >> lvl, w >> lvl)
        fl = 2.0 * ci * co * ks * ks * hh * ww
        fwd += fl
        if i == 0:
            first = fl
    return 3.0 * fwd - first if train else fwd


def tag_flops(tag: str, n: int) -> float:
    """Algorithmic FLOPs of one launch of a conv / wgrad kernel from its profiler tag 'name:k3:Cin>Cout@HxW'."""
    try:
        _, k, chans, = tag.split(":")[0:3]
        ks = int(k[1:])
        cc, hw = chans.split("@")
        ci, co = [int(v) for v in cc.split(">")]
        hh, ww = [int(v) for v in hw.split("x")]
        return 2.0 * ci * co * ks * ks * hh * ww * n
    except Exception:
        return 0.0


def tag_bytes(tag: str, n: int) -> float:
    """Algorithmic HBM bytes of one launch of a tagged elementwise kernel 'name:variant:C@HxW' (bf16 tensors, every tensor
    read or written once per pass)."""
    try:
        name, variant, rest = tag.split(":")[0:3]
        c, hw = rest.split("@")
        hh, ww = [int(v) for v in hw.split("x")]
        elems = float(n) * int(c) * hh * ww
        if name == "gn_bwd":                 # algorithmic minimum: read r and g once, write dr once (the kernel's second
            return 3 * 2.0 * elems           # phase re-reads both from HBM today: see roofline.traffic)
        if name == "gn_apply":               # read r, write a (+ pooled quarter)
            return (2 + (0.25 if variant.endswith("1") else 0.0)) * 2.0 * elems
        if name == "upsample2x_bwd":         # dims are the low-resolution result: read 4x gradient (+ addend), write 1x
            return (5 + (1 if variant.endswith("1") else 0)) * 2.0 * elems
    except Exception:
        pass
    return 0.0


def source_hash() -> str:
    """sha256 over the CUDA sources the library is built from: ties an ncu capture to the build it was taken from."""
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "mri_superresolution_b200", "csrc")
    for name in sorted(os.listdir(csrc)):
        if name.endswith((".cu", ".cuh", ".h")):
            h.update(name.encode())
            h.update(open(os.path.join(csrc, name), "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic():
    """DRAM bytes per step of each kernel family (dram__bytes_read.sum + dram__bytes_write.sum over the family's launches) from
    the newest committed ncu launch list profiles/rNN_traffic.json (tools/ncu_launch_list.py --json).  Only used when the capture
    was taken from THIS source tree (source_hash) and workload; otherwise roofline.traffic is null and says why."""
    prof = os.path.join(ROOT, "profiles")
    best = None
    for name in sorted(os.listdir(prof)) if os.path.isdir(prof) else []:
        if name.endswith("_traffic.json"):
            best = os.path.join(prof, name)
    if best is None:
        return None, "no ncu traffic capture committed"
    try:
        d = json.load(open(best))
    except Exception as exc:
        return None, f"{os.path.basename(best)} unreadable: {exc}"
    if d.get("source_hash") != source_hash():
        return None, (f"{os.path.basename(best)} was captured from source {d.get('source_hash')}, this build is "
                      f"{source_hash()}: stale, not reported")
    return d, "profiles/" + os.path.basename(best)


def peaks():
    p = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            d = json.load(open(path))
            for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained"):
                if k in d:
                    p[k] = float(d[k])
            p["source"] = "measured"
        except Exception:
            pass
    return p


# ----------------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ts, line in self.rows:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                clk = float(parts[1])
                smax = float(parts[2])
            except ValueError:
                continue
            if t0 <= ts <= t1 + 0.2:
                sm.append(clk)
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[4:8]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------- reference / CPU arm
CPU_BATCH = 4                   # BASELINE.json configs[0]: the reference's own CPU-runnable case is batch 4


def cpu_oracle_run(mode: str, steps: int, warmup: int, budget_s: float):
    """Times the oracle (reference algorithm, plain PyTorch fp32, all host threads) on a bounded sample of the
    workload: per step one batch of `b` slices of 256x256; b = 4 (BASELINE.json configs[0]: "batch 4, L1+SSIM, fp32 on CPU via
    the scripts/train.py path"), reduced only if the whole run would not fit `budget_s`."""
    import torch
    from oracle import unet_oracle as UO
    from oracle.synth import make_pairs
    from oracle.train_oracle import OracleTrainer, infer_step
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    params = UO.init_params(BASE_FILTERS, seed=SEED)
    trainer = OracleTrainer(params, lr=LR, weight_decay=WD, ssim_weight=SSIM_WEIGHT) if mode == "train" else None

    def one(b, seed):
        x, t = make_pairs(b, H, W, seed=seed)
        t0 = time.perf_counter()
        if mode == "train":
            trainer.step(x, t)
        else:
            infer_step(params, x)
        return time.perf_counter() - t0

    one(1, SEED)                                    # page-in / thread-pool warm-up, not counted
    t1 = one(1, SEED + 1)
    total = max(1, steps + warmup)
    b = int(max(1, min(CPU_BATCH, (budget_s / total) / max(t1, 1e-3))))
    for i in range(warmup):
        one(b, SEED + 10 + i)
    dt = 0.0
    for i in range(steps):
        dt += one(b, SEED + 100 + i)
    return {"value": b * steps / dt, "ms_per_step": 1e3 * dt / steps, "batch": b, "cores": cores,
            "threads": torch.get_num_threads()}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    mode = args.mode
    r = cpu_oracle_run(mode, args.steps, args.warmup, budget_s=150.0)
    unit = "slices/s"
    sample = (f"batch {r['batch']} slices of {H}x{W} per step x {args.steps} steps (BASELINE.json configs[0] batch size, the "
              f"bounded sample of the batch-{args.batch} workload), oracle ({'train step' if mode == 'train' else 'forward'}"
              f", fp32, torch CPU, {r['threads']} threads)")
    cfg = workload_config(mode, max(1, args.gpus), args.batch)
    cfg["reference_arm_batch_per_step"] = r["batch"]
    cfg["reference_arm_note"] = (f"the CPU arm runs batch {r['batch']} per step (configs[0] as written), not batch {args.batch}: "
                                 "slices/s is per-slice throughput, batch-size independent on the host within a few %")
    line = {
        "impl": "reference", "metric": METRIC if mode == "train" else INFER_METRIC, "value": r["value"], "unit": unit,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": r["value"], "unit": unit, "cores": r["threads"], "kind": "port", "sample": sample},
        "e2e": {"value": r["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(mode: str, world: int, per_gpu_batch: int, perceptual: float = 0.0):
    loss = "L1+SSIM(0.4)" if perceptual <= 0 else f"L1+SSIM(0.4)+VGG19 perceptual({perceptual})"
    cfg_id = (1 if perceptual <= 0 else 2) if mode == "train" else {512: 3, 320: 4}.get(H, 1)
    note = "" if mode == "train" or H == 256 else (" (one chunk of the batch-256 job)" if H == 512 else " (one rank's share of the 256-slice volume)")
    return {"workload": (f"BASELINE.json configs[{cfg_id}]{note}: U-Net(base_filters={BASE_FILTERS}) "
                         f"{'training step, ' + loss + ' loss, Adam' if mode == 'train' else 'inference forward'}, "
                         f"batch {per_gpu_batch}/GPU, {H}x{W} input -> {2 * H}x{2 * W} output"),
            "per_gpu_batch": per_gpu_batch, "global_batch": per_gpu_batch * world, "input": [H, W],
            "parallelism": f"dp{world}", "l2": "working set per step (>= 8 GB of activations) >> 126 MB L2; inputs re-read from HBM",
            **_non_default_build()}


def _non_default_build():
    """A/B runs are self-describing: the staged library variant and every MRISR_* kernel knob set in the environment (none in the
    driver's runs: the measured default build)."""
    knobs = {k: v for k, v in sorted(os.environ.items()) if k.startswith("MRISR_") and k != "MRISR_NONE"}
    return {"knobs": knobs} if knobs else {}


# ----------------------------------------------------------------------------------------- GPU eager baseline
def gpu_eager_run(dev, mode: str, nb: int, steps: int, warmup: int, amp: bool):
    """The reference algorithm in PyTorch eager on the SAME GPU (cuDNN / ATen kernels, one per op): oracle/ = the reference
    modules restated, fp32 with TF32 off, or under autocast(bf16) as scripts/train.py:304-311 runs with --use_amp.  This is the
    bar SURVEY.md names for a repository that ships no kernels of its own; a reported baseline, like cpu_baseline."""
    import torch
    from oracle import losses_oracle as LO
    from oracle import unet_oracle as UO
    from oracle.synth import make_pairs
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        params = {k: v.to(dev).requires_grad_(mode == "train") for k, v in UO.init_params(BASE_FILTERS, seed=SEED).items()}
        opt = torch.optim.Adam(list(params.values()), lr=LR, weight_decay=WD) if mode == "train" else None
        x, t = [v.to(dev) for v in make_pairs(nb, H, W, seed=SEED)]

        def one():
            if mode == "train":
                opt.zero_grad(set_to_none=True)
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                    out = UO.unet_forward(params, x)
                    loss = LO.combined_loss(out, t, ssim_weight=SSIM_WEIGHT)
                loss.backward()
                opt.step()
            else:
                with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                    torch.clamp(UO.unet_forward(params, x), 0, 1)

        for _ in range(warmup):
            one()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            one()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / steps
        peak_gb = torch.cuda.max_memory_allocated(dev) / 2 ** 30
        return {"value": nb / (ms * 1e-3), "unit": "slices/s", "ms_per_step": ms, "batch": nb, "steps": steps,
                "peak_memory_gib": round(peak_gb, 1)}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
        del params
        torch.cuda.empty_cache()


# ----------------------------------------------------------------------------------------- per-kernel attribution
def attribute(rep, prof_steps: int, nb: int, mode: str, ms_step: float):
    """Per-family table and roofline records from the live CUDA-event attribution (`rep` = [(tag, calls, total_ms)]).

    roofline            : the dominant family of the step (the contract's record)
    rooflines.loss_kernel : fused L1 + SSIM loss-and-gradient kernel against HBM (north_star: >= 70 %)
    rooflines.conv_layers : every convolution / weight-gradient family together against the sustained bf16 peak (north_star:
                            >= 60 %), algorithmic FLOPs = 2 Cin Cout k^2 H W per launch
    traffic             : DRAM bytes per launch from the committed ncu capture of THIS source tree, else null."""
    pk = peaks()
    traffic_db, traffic_src = ncu_traffic()
    fam = {}
    for tag, calls, tms in rep:
        name = tag.split(":")[0]
        f = fam.setdefault(name, {"ms": 0.0, "calls": 0, "flops": 0.0, "bytes": 0.0})
        f["ms"] += tms / prof_steps
        f["calls"] += calls // prof_steps
        f["flops"] += tag_flops(tag, nb) * calls / prof_steps
        f["bytes"] += tag_bytes(tag, nb) * calls / prof_steps
    if "l1_ssim" in fam:
        # fp32 pred + fp32 target read, fp32 gradient written (SURVEY.md 8(d) fp32 mode: 12 B per HR pixel); forward-only: 8 B
        fam["l1_ssim"]["bytes"] = (12.0 if mode == "train" else 8.0) * nb * (2 * H) * (2 * W) * fam["l1_ssim"]["calls"]
    total_ms = sum(f["ms"] for f in fam.values()) or 1.0
    families = []
    for name, f in sorted(fam.items(), key=lambda kv: -kv[1]["ms"]):
        families.append({"kernel": name, "ms_per_step": round(f["ms"], 4), "launches_per_step": f["calls"],
                         "share": round(f["ms"] / total_ms, 4),
                         "tflops": round(f["flops"] / (f["ms"] * 1e-3) / 1e12, 2) if f["flops"] else None,
                         "gbs": round(f["bytes"] / (f["ms"] * 1e-3) / 1e9, 1) if f["bytes"] and not f["flops"] else None})

    def traffic_of(name, calls):
        if traffic_db is None:
            return None, traffic_src
        if traffic_db.get("workload") != f"{mode}:{nb}x{H}x{W}":
            return None, f"{traffic_src} is a capture of {traffic_db.get('workload')}, not of this workload"
        v = traffic_db.get("families", {}).get(name)
        return (v / max(calls, 1), traffic_src) if v is not None else (None, f"{name} not in {traffic_src}")

    def record(name):
        f = fam[name]
        share = round(f["ms"] / total_ms, 4)
        base = {"kernel": name, "ms_per_step": round(f["ms"], 4), "launches_per_step": f["calls"], "share_of_step": share}
        tr, tr_src = traffic_of(name, f["calls"])
        if f["flops"] > 0:
            ach = f["flops"] / (f["ms"] * 1e-3) / 1e12
            return {**base, "bound": "tensor", "achieved": round(ach, 2), "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                    "frac": round(ach / pk["bf16_tflops_sustained"], 4), "traffic": tr, "traffic_source": tr_src,
                    "peak_source": pk["source"] + " (sustained bf16: kernel timed inside a long step)",
                    "flops_per_launch": f["flops"] / max(f["calls"], 1)}
        if f["bytes"] > 0:
            ach = f["bytes"] / (f["ms"] * 1e-3) / 1e9
            return {**base, "bound": "hbm", "achieved": round(ach, 1), "peak": pk["hbm_gbs"], "unit": "GB/s",
                    "frac": round(ach / pk["hbm_gbs"], 4), "traffic": tr, "traffic_source": tr_src, "peak_source": pk["source"],
                    "bytes_per_launch": f["bytes"] / max(f["calls"], 1),
                    "note": "algorithmic bytes = minimum passes x tensor size (DESIGN.md 4)"}
        return {**base, "bound": "hbm", "achieved": None, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": None, "traffic": tr,
                "traffic_source": tr_src, "note": "helper kernel without an algorithmic-byte model; its time counts against the step"}

    roofline = record(families[0]["kernel"])
    rooflines = {}
    if "l1_ssim" in fam:
        rooflines["loss_kernel"] = record("l1_ssim")
    conv = [(k, v) for k, v in fam.items() if v["flops"] > 0]
    if conv:
        fl, t = sum(v["flops"] for _, v in conv), sum(v["ms"] for _, v in conv)
        ach = fl / (t * 1e-3) / 1e12
        rooflines["conv_layers"] = {
            "kernels": sorted(k for k, _ in conv), "bound": "tensor", "achieved": round(ach, 2),
            "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s", "frac": round(ach / pk["bf16_tflops_sustained"], 4),
            "ms_per_step": round(t, 4), "share_of_step": round(t / total_ms, 4), "flops_per_step": fl,
            "per_family": {k: {"tflops": round(v["flops"] / (v["ms"] * 1e-3) / 1e12, 1), "ms": round(v["ms"], 4),
                               "frac": round(v["flops"] / (v["ms"] * 1e-3) / 1e12 / pk["bf16_tflops_sustained"], 4)}
                           for k, v in sorted(conv, key=lambda kv: -kv[1]["ms"])}}
    flops_step = unet_flops_per_slice(BASE_FILTERS, H, W, mode == "train") * nb
    whole = {"unet_conv_tflops": round(flops_step / (ms_step * 1e-3) / 1e12, 2),
             "frac_of_bf16_sustained": round(flops_step / (ms_step * 1e-3) / 1e12 / pk["bf16_tflops_sustained"], 4),
             "attributed_ms": round(total_ms, 4)}
    return families, roofline, rooflines, whole


# ----------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from mri_superresolution_b200 import _native
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    from mri_superresolution_b200.optim import FusedAdam
    from mri_superresolution_b200.parallel import GradientBuckets
    from mri_superresolution_b200.utils.losses import CombinedLoss
    from mri_superresolution_b200.utils.synthetic import make_pairs

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    mode, nb = args.mode, args.batch
    _native.lib()                                   # fail loudly if the CUDA library is missing

    torch.manual_seed(SEED)
    model = UNetSuperRes(base_filters=BASE_FILTERS, precision=args.precision).to(dev)
    if args.perceptual > 0:
        from mri_superresolution_b200.utils.synthetic import make_vgg19_weights
        crit = CombinedLoss(ssim_weight=SSIM_WEIGHT, perceptual_weight=args.perceptual, vgg_weights=make_vgg19_weights())
    else:
        crit = CombinedLoss(ssim_weight=SSIM_WEIGHT)
    opt = FusedAdam(model, lr=LR, weight_decay=WD, capturable=True)
    buckets = GradientBuckets(model) if world > 1 else None
    x_cpu, t_cpu = make_pairs(nb, H, W, seed=SEED, rank=rank)
    x_pin, t_pin = x_cpu.pin_memory(), t_cpu.pin_memory()
    x_dev, t_dev = x_cpu.to(dev), t_cpu.to(dev)

    def train_step(x, t):
        opt.zero_grad(set_to_none=True)
        out = model(x)
        loss = crit(out, t)
        loss.backward()
        scale = buckets.allreduce() if buckets is not None else 1.0
        opt.step(grad_scale=scale)
        return loss

    def infer_step(x, t):
        with torch.no_grad():
            return torch.clamp(model(x), 0, 1)

    step = train_step if mode == "train" else infer_step
    if mode == "infer":
        model.eval()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k, tail=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        if tail is not None:
            tail()                               # work still in flight on another stream joins before the closing event
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- resident-input timing -------------------------------------------------------------
    # the training step is replayed as one CUDA graph (utils/graph.py: the public GraphedTrainStep), so the GPU never waits
    # for the ~190 launches of a step to be enqueued; under data parallelism the bucketed NCCL all-reduces on the side stream
    # are part of the graph.  gpu_launches counts the kernels one step launches (host counter over an eager step).
    for _ in range(max(args.warmup - 1, 1)):
        step(x_dev, t_dev)
    l0 = _native.launch_count()
    step(x_dev, t_dev)
    launches_per_step = _native.launch_count() - l0
    graphed = None
    graphed_obj = None
    graph_note = None
    try:
        if mode == "train" and not args.no_graph:
            from mri_superresolution_b200.utils.graph import GraphedTrainStep
            graphed = graphed_obj = GraphedTrainStep(model, crit, opt, x_dev, t_dev, buckets=buckets)
        elif mode == "infer" and not args.no_graph:
            from mri_superresolution_b200.utils.graph import GraphedForward
            graphed_obj = gfwd = GraphedForward(model, x_dev)
            graphed = lambda x, t: gfwd(x)                               # noqa: E731
    except Exception as exc:             # same kernels either way: a failed capture only changes how they are launched
        if world > 1:
            raise                        # ranks must not diverge (one replaying, one launching eagerly would still match, but say so loudly)
        graphed, graphed_obj, graph_note = None, None, f"graph capture failed, eager launches: {type(exc).__name__}: {exc}"
        print("bench.py: " + graph_note, file=sys.stderr, flush=True)
        torch.cuda.synchronize()
    run_step = (lambda x, t: graphed(x, t)) if graphed is not None else step
    run_step(x_dev, t_dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    t_wall0 = time.time()
    ms = timed(lambda: run_step(x_dev, t_dev), args.steps)
    t_wall1 = time.time()
    launches = launches_per_step * args.steps
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    value = nb * world * args.steps / (ms * 1e-3)

    # ---- end to end: pinned host inputs, H2D copies and loss read-back inside the timed region
    # Every step's batch is copied host -> device once, by the package's double-buffered DevicePrefetcher (the copy of batch
    # i+1 runs on a side stream while step i computes; utils/prefetch.py), as a training loop built on this package would.
    from mri_superresolution_b200.utils.prefetch import DevicePrefetcher
    d2h = [0]
    out_pin = [None]
    e2e_steps = max(2, min(args.steps, 10))

    def host_batches():
        while True:
            yield (x_pin, t_pin) if mode == "train" else (x_pin,)         # inference needs no target

    feeder = DevicePrefetcher(host_batches(), dev)
    def e2e_step():
        batch = next(feeder)
        if graphed is not None:
            r = graphed(batch[0], batch[1] if mode == "train" else None)
        else:
            r = step(batch[0], batch[1] if mode == "train" else None)
        if mode == "train":
            r.item()                                 # scripts/train.py:321 reads the loss every step
            d2h[0] = 4
        else:
            if out_pin[0] is None:
                out_pin[0] = torch.empty(r.shape, dtype=r.dtype).pin_memory()
            out_pin[0].copy_(r, non_blocking=True)   # the step's result is the whole super-resolved batch
            torch.cuda.current_stream().synchronize()
            d2h[0] = r.numel() * r.element_size()

    e2e_step()
    bytes0 = feeder.bytes_copied
    ms_e2e = timed(e2e_step, e2e_steps)
    e2e_value = nb * world * e2e_steps / (ms_e2e * 1e-3)
    h2d = (feeder.bytes_copied - bytes0) // e2e_steps          # counted from the tensors copied inside the timed region

    # ---- per-kernel attribution (same workload, CUDA events on the launching stream) ---------
    roofline, rooflines, families, whole = None, None, [], None
    prof_steps = 2
    barrier()
    if rank == 0:
        _native.profile(True)
    for _ in range(prof_steps):                      # every rank runs the steps (they contain the gradient all-reduce)
        step(x_dev, t_dev)
    barrier()
    if rank == 0:
        rep = _native.profile_report()
        _native.profile(False)
        if args.layer_report:
            with open(args.layer_report, "w") as fh:
                for tag, calls, tms in sorted(rep, key=lambda r: -r[2]):
                    fl = tag_flops(tag, nb) * calls / prof_steps
                    ms_t = tms / prof_steps
                    fh.write("%-44s calls/step %3d  ms/step %8.4f  TFLOP/s %s\n" % (
                        tag, calls // prof_steps, ms_t, ("%.1f" % (fl / (ms_t * 1e-3) / 1e12)) if fl else "-"))
        families, roofline, rooflines, whole = attribute(rep, prof_steps, nb, mode, ms / args.steps)

    # ---- the metric is "train & infer": the inference forward of the same model / batch, inside the same line --------------
    infer = None
    if mode == "train" and not args.no_infer:
        from mri_superresolution_b200.utils.graph import GraphedForward
        model.eval()
        gf = GraphedForward(model, x_dev)
        l0 = _native.launch_count()
        with torch.no_grad():
            model(x_dev)
        infer_launches = _native.launch_count() - l0
        for _ in range(3):
            gf(x_dev)
        ms_i = timed(lambda: gf(x_dev), args.steps)
        feeder_i = DevicePrefetcher(((x_pin,) for _ in iter(int, 1)), dev)
        y_pin = torch.empty((nb, 1, 2 * H, 2 * W), dtype=torch.float32).pin_memory()

        def infer_e2e():
            (xb,) = next(feeder_i)
            y_pin.copy_(gf(xb), non_blocking=True)
            torch.cuda.current_stream().synchronize()

        infer_e2e()
        b0 = feeder_i.bytes_copied
        ms_ie = timed(infer_e2e, e2e_steps)
        infer = {"metric": INFER_METRIC, "value": nb * world * args.steps / (ms_i * 1e-3), "unit": "slices/s",
                 "ms_per_step": ms_i / args.steps, "steps": args.steps, "gpu_launches": infer_launches * args.steps,
                 "e2e": {"value": nb * world * e2e_steps / (ms_ie * 1e-3), "unit": "slices/s",
                         "h2d_bytes_per_step": (feeder_i.bytes_copied - b0) // e2e_steps,
                         "d2h_bytes_per_step": y_pin.numel() * 4, "ms_per_step": ms_ie / e2e_steps,
                         "pipeline": "pinned host batch -> DevicePrefetcher -> GraphedForward -> whole output batch copied to "
                                     "pinned host memory every step"},
                 "workload": f"inference forward (scripts/infer.py:268-276) of the same model, batch {nb}/GPU, {H}x{W} -> {2 * H}x{2 * W}",
                 "unet_conv_tflops": round(unet_flops_per_slice(BASE_FILTERS, H, W, False) * nb / (ms_i / args.steps * 1e-3) / 1e12, 2)}
        # the same forward in the formats scripts/infer.py moves: 8-bit gray levels in (PNG, :97), percentile window + min-max
        # on the device (:97-130), 8-bit images out ((x * 255).astype(uint8), :331); the copy of step i's images overlaps step i + 1
        from mri_superresolution_b200.utils.preprocess import preprocess_images
        u8_pin = (torch.rand(nb, H, W, generator=torch.Generator().manual_seed(SEED)) * 255).to(torch.uint8).pin_memory()
        y8_pin = [torch.empty((nb, 2 * H, 2 * W), dtype=torch.uint8).pin_memory() for _ in range(2)]
        y8_dev = [None, None]
        ev_ready = [torch.cuda.Event() for _ in range(2)]
        ev_done = [torch.cuda.Event() for _ in range(2)]
        copy_stream = torch.cuda.Stream()
        slot = [0]

        def infer_png():
            k = slot[0]
            y = gf(preprocess_images(u8_pin.to(dev, non_blocking=True)))
            y8_dev[k] = (y[:, 0] * 255).to(torch.uint8)
            ev_ready[k].record()
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ev_ready[k])
                y8_pin[k].copy_(y8_dev[k], non_blocking=True)
                ev_done[k].record()
            slot[0] = k ^ 1
            ev_done[k ^ 1].synchronize()                  # the previous step's images are on the host

        with torch.cuda.stream(copy_stream):
            for e in ev_done:
                e.record()
        for _ in range(3):
            infer_png()
        ms_ip = timed(infer_png, e2e_steps, tail=lambda: torch.cuda.current_stream().wait_stream(copy_stream))
        copy_stream.synchronize()
        infer["e2e_png_formats"] = {
            "value": nb * world * e2e_steps / (ms_ip * 1e-3), "unit": "slices/s", "ms_per_step": ms_ip / e2e_steps,
            "h2d_bytes_per_step": nb * H * W, "d2h_bytes_per_step": nb * 4 * H * W,
            "pipeline": "pinned uint8 slices -> device -> percentile window + min-max (mrisr_percentile_normalize_u8) -> "
                        "GraphedForward -> x 255 truncated to uint8 -> pinned host, the copy overlapping the next step"}
        gf.close()
        del gf, feeder_i
        model.train()

    # ---- the reference algorithm in PyTorch eager on the same GPU (rank 0, N = 1 only) ------------------------------------
    eager = None
    if rank == 0 and world == 1 and not args.no_eager_baseline:
        eager = {"what": "oracle/ (= the reference modules' algorithm) in PyTorch eager on this GPU: cuDNN/ATen kernels, one per "
                         "op, same batch; a reported baseline (SURVEY.md: the reference ships no kernels of its own)"}
        for key, amp in (("fp32_tf32_off", False), ("autocast_bf16", True)):
            try:
                eager[key] = gpu_eager_run(dev, mode, nb, steps=5, warmup=3, amp=amp)
            except Exception as exc:                      # e.g. out of memory beside the resident workspaces
                eager[key] = {"unavailable": f"{type(exc).__name__}: {str(exc)[:200]}"}
                torch.cuda.empty_cache()

    # ---- CPU baseline (rank 0, N = 1 only) ----------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_oracle_run(mode, steps=4, warmup=1, budget_s=25.0)
        cpu = {"value": r["value"], "unit": "slices/s", "cores": r["threads"], "kind": "port",
               "sample": f"batch {r['batch']} slices of {H}x{W} per step x 4 steps of the oracle "
                         f"({'train step' if mode == 'train' else 'forward'}, fp32, torch CPU; BASELINE.json configs[0] batch size)"}

    if rank == 0:
        line = {
            "metric": METRIC if mode == "train" else INFER_METRIC, "value": value, "unit": "slices/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": workload_config(mode, world, nb, args.perceptual),
            "e2e": {"value": e2e_value, "unit": "slices/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h[0],
                    "steps": e2e_steps, "ms_per_step": ms_e2e / e2e_steps,
                    "pipeline": "pinned host batch -> DevicePrefetcher (one H2D copy per step, on a side stream, double-buffered) "
                                "-> " + (("GraphedTrainStep" if mode == "train" else "GraphedForward") + " (whole step replayed as one CUDA graph)" if graphed is not None else "public modules (eager launches)")
                                + " -> result read back every step"},
            "gpu_launches": launches, "launch_mode": ("cuda graph replay of the eagerly counted step" if graphed is not None else (graph_note or "eager")),
            "clocks": clocks, "roofline": roofline, "rooflines": rooflines, "cpu_baseline": cpu,
            "infer": infer, "gpu_eager_baseline": eager,
            "kernels": families, "whole_step": whole,
        }
        print(json.dumps(line), flush=True)
    sys.stdout.flush()
    # Teardown order matters under data parallelism: the recorded step holds captured NCCL all-reduces, and NCCL's communicator
    # teardown waits for every graph that references it.  Destroy the graph, the prefetcher and the buckets first; the JSON line
    # is already out, so a watchdog ends the process if the communicator still refuses to go (round 1: rc 124 at N = 2).
    if hasattr(graphed_obj, "close"):
        graphed_obj.close()
    del graphed, graphed_obj, run_step, feeder
    if buckets is not None:
        buckets.close()
    import gc
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        watchdog = threading.Timer(20.0, lambda: os._exit(0))
        watchdog.daemon = True
        watchdog.start()
        dist.destroy_process_group()
        watchdog.cancel()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="train", choices=["train", "infer"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--batch", type=int, default=PER_GPU_BATCH, help="slices per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-infer", action="store_true", help="skip the inference sub-record of a training run")
    ap.add_argument("--no-eager-baseline", action="store_true", help="skip the PyTorch-eager-on-GPU baseline leg")
    ap.add_argument("--no-graph", action="store_true", help="end-to-end leg with eager launches instead of the CUDA graph")
    ap.add_argument("--perceptual", type=float, default=0.0,
                    help="weight of the VGG19[:36] perceptual term (BASELINE.json configs[2] uses 0.1); 0 = configs[1]")
    ap.add_argument("--input", type=int, default=256,
                    help="square network input size; 256 = the headline workload, 512 / 320 = the slice sizes of BASELINE.json "
                         "configs 4 / 5 (use with --mode infer and a --batch that is one rank's chunk)")
    ap.add_argument("--layer-report", default=None, help="write the per-kernel-tag (per layer shape) timing table here")
    ap.add_argument("--base-filters", type=int, default=32,
                    help="32 = the headline workload (scripts/train.py:501); 64 = the default of the reference's scripts/infer.py:473")
    args = ap.parse_args()
    global H, W, BASE_FILTERS
    H = W = args.input
    BASE_FILTERS = args.base_filters
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and world == 1 and args.gpus > 1:
        # convenience: re-launch under torch.distributed.run when called as `python bench.py --gpus N`
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_ours(args)


if __name__ == "__main__":
    main()
