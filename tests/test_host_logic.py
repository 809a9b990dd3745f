"""CPU-side tests: the C-ABI library loads and exports every symbol include/mrisr.h declares, argument
validation mirrors the reference, checkpoint layout matches, and the data-parallel bucket logic
(world_size 2, gloo).  No compute calls (there is no GPU here)."""
import os
import re
import socket

import numpy as np

import pytest
import torch
import torch.multiprocessing as mp

from oracle import unet_oracle as UO

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from mri_superresolution_b200 import _native
    L = _native.lib()
    hdr = open(os.path.join(ROOT, "include", "mrisr.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(mrisr_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 13
    for name in declared:
        assert hasattr(L, name), name
    assert declared == set(_native.SIGNATURES), declared ^ set(_native.SIGNATURES)
    assert L.mrisr_version() == 100


def test_param_arena_layout_matches_reference_order():
    from mri_superresolution_b200 import _native
    from mri_superresolution_b200.models.unet_model import param_specs
    L = _native.lib()
    for f in (16, 32, 64):
        specs = param_specs(f)
        assert specs == UO.param_specs(f)
        assert L.mrisr_unet_param_tensors(f) == len(specs) == 64
        off = 0
        for i, (_, shape) in enumerate(specs):
            n = 1
            for s in shape:
                n *= s
            assert L.mrisr_unet_param_offset(f, i) == off
            assert L.mrisr_unet_param_size(f, i) == n
            off += n
        assert L.mrisr_unet_param_elems(f) == off
    assert L.mrisr_unet_param_elems(32) == 1823122


def test_shape_validation_and_workspace_query():
    from mri_superresolution_b200 import _native
    L = _native.lib()
    assert L.mrisr_unet_workspace_bytes(2, 32, 32, 16, _native.FP32, 1) > L.mrisr_unet_workspace_bytes(2, 32, 32, 16, _native.FP32, 0) > 0
    assert L.mrisr_unet_workspace_bytes(2, 30, 32, 16, _native.FP32, 1) == 0        # H % 8 != 0
    assert b"multiples of 8" in L.mrisr_last_error()
    assert L.mrisr_unet_workspace_bytes(2, 32, 32, 24, _native.FP32, 1) == 0        # f % 16 != 0
    assert L.mrisr_unet_workspace_bytes(2, 32, 32, 16, 7, 1) == 0                   # unknown precision
    # the smallest legal slices (one pixel at the deepest level): the size arithmetic used to divide by the zero pooled width
    for h, w in ((8, 8), (16, 8), (8, 24)):
        for f in (16, 32, 64):
            for prec in (_native.BF16, _native.FP32):
                assert L.mrisr_unet_workspace_bytes(1, h, w, f, prec, 1) > L.mrisr_unet_workspace_bytes(1, h, w, f, prec, 0) > 0
    rc = L.mrisr_unet_forward(None, None, None, None, 0, 1, 8, 8, 16, 0, 0, None)
    assert rc == -1


def test_module_mirrors_reference_interface():
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    m = UNetSuperRes(in_channels=1, out_channels=1, base_filters=32, initial_alpha=50.0)
    assert (m.in_channels, m.out_channels, m.base_filters) == (1, 1, 32)
    assert abs(m.alpha.item() - 0.5) < 1e-7
    sd = m.state_dict()
    assert list(sd.keys()) == [n for n, _ in UO.param_specs(32)]
    assert all(tuple(sd[n].shape) == s for n, s in UO.param_specs(32))
    assert [n for n, _ in m.named_parameters()] == list(sd.keys())
    m.load_state_dict(UO.init_params(32), strict=True)      # strict load of a reference-style state dict
    assert m._is_flat()
    with pytest.raises(RuntimeError):
        m.load_state_dict({"alpha": torch.tensor(0.0)}, strict=True)
    with pytest.raises(ValueError):
        UNetSuperRes(in_channels=3)
    with pytest.raises(ValueError):
        UNetSuperRes(base_filters=24)
    with pytest.raises(RuntimeError):                  # no CPU fallback
        m(torch.zeros(1, 1, 16, 16))
    # kaiming fan_out statistics of the initialiser (models/unet_model.py:181)
    w = UNetSuperRes(base_filters=32).get_parameter("down3.maxpool_conv.1.double_conv.3.weight")
    assert abs(w.std().item() - (2.0 / (1 + 0.01 ** 2) / (256 * 9)) ** 0.5) < 2e-3


def test_checkpoint_roundtrip(tmp_path):
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    a, b = UNetSuperRes(base_filters=16), UNetSuperRes(base_filters=16)
    path = tmp_path / "best_model_unet.pth"
    torch.save({"epoch": 3, "model_state_dict": a.state_dict(), "val_loss": 0.1, "val_ssim": 0.9}, path)   # scripts/train.py:411-418
    ck = torch.load(path, map_location="cpu")
    b.load_state_dict(ck["model_state_dict"])                                                             # scripts/infer.py:59-61
    assert torch.equal(a.flat_parameters(), b.flat_parameters())


def test_loss_module_validation_and_attributes():
    from mri_superresolution_b200.utils.losses import CombinedLoss, SSIM, create_window, gaussian_window
    for sw, pw in ((-0.1, 0.0), (1.1, 0.0), (0.5, -0.1), (0.7, 0.5)):
        with pytest.raises(ValueError):
            CombinedLoss(ssim_weight=sw, perceptual_weight=pw)
    c = CombinedLoss(ssim_weight=0.4)
    assert abs(c.l1_weight - 0.6) < 1e-12 and c.use_perceptual is False and c.perceptual_loss is None
    assert c.window.shape == (1, 1, 11, 11) and (c.window_size, c.sigma, c.val_range) == (11, 1.5, 1.0)
    from oracle.losses_oracle import gaussian_taps, gaussian_window2d
    assert torch.equal(gaussian_window(11, 1.5), gaussian_taps())
    assert torch.allclose(create_window(11, 1, 1.5, "cpu")[0, 0], gaussian_window2d(), atol=0, rtol=0)
    assert SSIM(device=torch.device("cpu")).window.shape == (1, 1, 11, 11)
    with pytest.raises(RuntimeError):
        c(torch.zeros(1, 1, 16, 16), torch.zeros(1, 1, 16, 16))     # CPU tensors: no fallback


def test_bucket_bounds_cover_arena():
    from mri_superresolution_b200.parallel import bucket_bounds
    for total, nb in ((1823122, 3), (10, 3), (7, 10), (5, 1)):
        b = bucket_bounds(total, nb)
        assert sorted(b)[0][0] == 0 and sorted(b)[-1][1] == total
        assert sum(e - s for s, e in b) == total
        assert b[0][1] == total                      # last slice of the arena goes first


def test_schedule_buckets_follow_the_backward_order():
    """Gradient-readiness marks (include/mrisr.h): the offset from which gradients are final never grows along the backward
    schedule 19, 17, 18, 16, ..., 0, ends at 0, and the schedule-aligned buckets tile the arena back to front."""
    from mri_superresolution_b200 import _native
    from mri_superresolution_b200.parallel import SCHEDULE_CUTS, schedule_bounds
    L = _native.lib()
    for f in (16, 32):
        total = L.mrisr_unet_param_elems(f)
        order = [19, 17, 18] + list(range(16, -1, -1))
        ff = [L.mrisr_unet_grads_final_from(f, l) for l in order]
        assert all(a >= b for a, b in zip(ff, ff[1:])) and ff[-1] == 0 and 0 < ff[0] < total
        # a block's own weight tensor is final once the block is done
        assert L.mrisr_unet_grads_final_from(f, 9) == L.mrisr_unet_param_offset(f, 28)
        b = schedule_bounds(total, [L.mrisr_unet_grads_final_from(f, l) for l in SCHEDULE_CUTS])
        assert b[0][1] == total and b[-1][0] == 0 and all(x[0] == y[1] for x, y in zip(b, b[1:]))
        assert (b[-1][1] - b[-1][0]) * 4 < 300e3 * (f / 32) ** 2      # the bucket exchanged after backward ends is small
    assert L.mrisr_unet_grads_final_from(32, 20) == -1
    assert schedule_bounds(10, [7, 7, 3]) == [(7, 10), (3, 7), (0, 3)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _dp_worker(rank, world, port, q):
    import torch.distributed as dist
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    from mri_superresolution_b200.parallel import GradientBuckets
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    torch.manual_seed(0)
    model = UNetSuperRes(base_filters=16)
    # emulate an arena-backed backward on CPU: rank-dependent gradients written as arena views
    arena = torch.empty(model._total)
    model._grad_arenas.append(arena)
    g = torch.Generator().manual_seed(100 + rank)
    arena.copy_(torch.randn(model._total, generator=g))
    for p, o in zip(model._param_list(), model._offs):
        p.grad = arena[o:o + p.numel()].view(p.shape)
    scale = GradientBuckets(model, n_buckets=3).allreduce()
    expect = sum(torch.randn(model._total, generator=torch.Generator().manual_seed(100 + r)) for r in range(world))
    ok = torch.allclose(model.flat_grads(), expect, atol=1e-6) and abs(scale - 1.0 / world) < 1e-12
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_data_parallel_bucketed_allreduce_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_perceptual_argument_validation_mirrors_reference():
    """utils/losses.py:133 raises ValueError for an unknown perceptual loss type; the mirror also refuses feature stacks
    that do not end on conv + ReLU (the reference default, idx 35 = relu5_4, does)."""
    from oracle import losses_oracle as LO
    from mri_superresolution_b200.utils.perceptual import PerceptualLoss, VGGFeatureExtractor
    vggp = LO.make_vgg_params(seed=1235)
    with pytest.raises(ValueError):
        PerceptualLoss(35, "huber", vgg_weights=vggp)
    with pytest.raises(ValueError):
        VGGFeatureExtractor(36, weights=vggp)                     # ends on the last MaxPool
    ext = VGGFeatureExtractor(35, weights=vggp)
    assert ext.cfg == [64, 64, -1, 128, 128, -1, 256, 256, 256, 256, -1, 512, 512, 512, 512, -1, 512, 512, 512, 512]
    assert len(ext.features) == 36 and not any(p.requires_grad for p in ext.features.parameters())
    assert tuple(ext.mean.shape) == (1, 3, 1, 1) and tuple(ext.std.shape) == (1, 3, 1, 1)
    with pytest.raises(RuntimeError):
        ext(torch.zeros(1, 1, 32, 32))                            # CPU tensor: no fallback


def test_vgg_abi_size_queries():
    import ctypes
    from mri_superresolution_b200 import _native
    L = _native.lib()
    cfg = [64, 64, -1, 128, 128, -1, 256, 256, 256, 256, -1, 512, 512, 512, 512, -1, 512, 512, 512, 512]
    arr = (ctypes.c_int * len(cfg))(*cfg)
    n_w = sum(co * ci * 9 + co for ci, co in zip([3] + [c for c in cfg if c > 0][:-1], [c for c in cfg if c > 0]))
    assert L.mrisr_vgg_weight_elems(arr, len(cfg)) == n_w
    assert L.mrisr_vgg_packed_bytes(arr, len(cfg)) > 2 * 2 * (n_w - sum(c for c in cfg if c > 0)) - 1
    assert L.mrisr_vgg_workspace_bytes(2, 64, 64, arr, len(cfg), 1) > L.mrisr_vgg_workspace_bytes(2, 64, 64, arr, len(cfg), 0) > 0
    assert L.mrisr_vgg_workspace_bytes(2, 40, 64, arr, len(cfg), 1) == 0             # H % 16 != 0
    bad = (ctypes.c_int * 2)(64, -1)
    assert L.mrisr_vgg_packed_bytes(bad, 2) == 0                                      # must end on a conv


def test_device_prefetcher_rejects_cpu_device():
    from mri_superresolution_b200.utils.prefetch import DevicePrefetcher
    with pytest.raises(RuntimeError):
        DevicePrefetcher([], "cpu")


# ------------------------------------------------------------------ drivers and optimizer state layout
def test_log_message_schema_matches_reference(capsys):
    """scripts/train.py:54-68: dict messages get floats rounded to 6 places and a 'type' key; anything else is a string message."""
    import json
    from mri_superresolution_b200.drivers import log_message
    log_message({"epoch": 3, "train_loss": 0.123456789, "val_loss": "N/A"}, "epoch_summary")
    log_message(RuntimeError("boom"))
    lines = [json.loads(l) for l in capsys.readouterr().out.strip().splitlines()]
    assert lines[0] == {"epoch": 3, "train_loss": 0.123457, "val_loss": "N/A", "type": "epoch_summary"}
    assert lines[1] == {"type": "info", "message": "boom"}


def test_find_best_checkpoint_order_and_errors(tmp_path):
    """scripts/infer.py:74-95: best_ before final_ before any *.pth naming the model type; FileNotFoundError otherwise."""
    from mri_superresolution_b200.drivers import TrainConfig, find_best_checkpoint
    with pytest.raises(FileNotFoundError):
        find_best_checkpoint(str(tmp_path), "unet")
    (tmp_path / "epoch3_unet.pth").write_bytes(b"x")
    assert find_best_checkpoint(str(tmp_path), "unet").endswith("epoch3_unet.pth")
    (tmp_path / "final_model_unet.pth").write_bytes(b"x")
    assert find_best_checkpoint(str(tmp_path), "unet").endswith("final_model_unet.pth")
    (tmp_path / "best_model_unet.pth").write_bytes(b"x")
    assert find_best_checkpoint(str(tmp_path), "unet").endswith("best_model_unet.pth")
    with pytest.raises(ValueError):
        TrainConfig(model_type="resnet")                                   # scripts/train.py:175


def test_fused_adam_state_dict_is_stock_adam_layout():
    """optimizer_state_dict of a checkpoint (scripts/train.py:411-418) is torch.optim.Adam's: it must load into FusedAdam and
    FusedAdam's must load into torch.optim.Adam over the same parameters, values preserved."""
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    from mri_superresolution_b200.optim import FusedAdam
    torch.manual_seed(0)
    model = UNetSuperRes(base_filters=16)
    ref_params = [torch.nn.Parameter(p.detach().clone()) for p in model.parameters()]
    ref = torch.optim.Adam(ref_params, lr=3e-4, weight_decay=1e-5)
    for p in ref_params:
        p.grad = torch.randn_like(p)
    ref.step()
    ref.step()
    sd = ref.state_dict()
    opt = FusedAdam(model, lr=1e-4)
    assert opt.state_dict()["state"] == {}                                 # torch: empty before the first step
    opt.load_state_dict(sd)
    assert opt.step_count == 2 and opt.lr == 3e-4 and opt.weight_decay == 1e-5
    back = opt.state_dict()
    assert sorted(back["state"].keys()) == sorted(sd["state"].keys())
    for i, st in sd["state"].items():
        assert torch.equal(back["state"][i]["exp_avg"], st["exp_avg"]) and torch.equal(back["state"][i]["exp_avg_sq"], st["exp_avg_sq"])
        assert float(back["state"][i]["step"]) == float(st["step"])
    assert back["param_groups"][0]["params"] == sd["param_groups"][0]["params"]
    fresh = torch.optim.Adam([torch.nn.Parameter(p.detach().clone()) for p in model.parameters()], lr=1.0)
    fresh.load_state_dict(back)                                            # stock Adam accepts ours
    assert fresh.param_groups[0]["lr"] == 3e-4
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.5, patience=0)      # scripts/train.py:189
    sched.step(1.0)
    sched.step(2.0)
    assert opt.lr == 1.5e-4


def test_gn_bwd_incremental_addressing_replay(tmp_path):
    """csrc/gn_bwd_trk.h (incremental addressing of gn_bwd_fused_kernel, plain-tensor kinds): tests/native/gn_bwd_trk_replay.cpp walks every
    (cluster rank, thread) through the kernel's issue / reload / apply order with the constant-step tracker and compares every
    operand index with the direct padded-pixel form of the default build; each unit must be issued and applied exactly once."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "trk_replay")
    subprocess.run(["g++", "-O1", "-std=c++17", "-I", os.path.join(root, "mri_superresolution_b200", "csrc"),
                    os.path.join(root, "tests", "native", "gn_bwd_trk_replay.cpp"), "-o", exe], check=True)
    res = subprocess.run([exe], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]
    assert res.stdout.startswith("ok ")


def _run_host_kernel_body(tmp_path, name):
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / name)
    subprocess.run([nvcc, "-std=c++17", "-O1", "--expt-relaxed-constexpr", "-Wno-deprecated-gpu-targets", "-I",
                    os.path.join(root, "mri_superresolution_b200", "csrc"),
                    os.path.join(root, "tests", "native", name + ".cu"), "-o", exe], check=True)
    res = subprocess.run([exe], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]
    assert res.stdout.startswith("ok ")


def test_upsample_slide_body_runs_on_host(tmp_path):
    """csrc/upsample_slide.cuh (the bilinear x2 forward kernel): the kernel body is a __host__ __device__ function of
    (sample, block, thread); tests/native/upsample_slide_host.cu executes every thread of the launch grid on the CPU (NHWC and
    planar-8, fp32 and bf16, odd and one-row shapes) and compares bit for bit with the default kernel's direct form."""
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "upsample_slide_host")
    subprocess.run([nvcc, "-std=c++17", "-O1", "--expt-relaxed-constexpr", "-Wno-deprecated-gpu-targets", "-I",
                    os.path.join(root, "mri_superresolution_b200", "csrc"),
                    os.path.join(root, "tests", "native", "upsample_slide_host.cu"), "-o", exe], check=True)
    res = subprocess.run([exe], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]
    assert res.stdout.startswith("ok ")


def test_missing_library_fails_loudly(tmp_path):
    """_native.lib(): no library -> RuntimeError naming the build command (no silent CPU or PyTorch fallback)."""
    from mri_superresolution_b200 import _native
    saved, saved_lib = _native.LIB_PATH, _native._lib
    try:
        _native.LIB_PATH, _native._lib = str(tmp_path / "libmrisr_b200.so"), None
        with pytest.raises(RuntimeError, match="no CPU"):
            _native.lib()
    finally:
        _native.LIB_PATH, _native._lib = saved, saved_lib


def test_workspace_cache_never_drops_a_pinned_entry():
    """utils/workspace.py: a recorded CUDA graph holds raw pointers into its workspace, so pinned entries survive any number of
    other shapes (a ragged last batch or a no-grad validation pass used to evict the graphed workspace)."""
    import torch
    from mri_superresolution_b200.utils.workspace import WorkspaceCache
    made = []

    def alloc(tag):
        def f():
            made.append(tag)
            return torch.zeros(4)
        return f

    c = WorkspaceCache(keep=2)
    a = c.get(("a",), True, alloc("a"))
    assert c.get(("a",), True, alloc("a2")) is a and made == ["a"]
    token = c.pin_all()
    for i in range(5):                                   # five other train-kind shapes come and go
        c.get(("t", i), True, alloc(f"t{i}"))
    assert ("a",) in c and c.get(("a",), True, alloc("a3")) is a
    assert sum(1 for k in c.keys() if k[0] == "t") == 2  # the two most recent un-pinned entries are kept, no more
    assert ("t", 4) in c and ("t", 3) in c
    e = c.get(("e",), False, alloc("e"))                 # another kind does not compete with the train kind
    assert ("t", 4) in c and c.get(("e",), False, alloc("e2")) is e
    # alternating between two un-pinned shapes (full and ragged batch) does not re-allocate
    n = len(made)
    for _ in range(3):
        c.get(("t", 4), True, alloc("x"))
        c.get(("t", 3), True, alloc("y"))
    assert len(made) == n
    c.unpin(token)
    for i in range(5, 8):
        c.get(("t", i), True, alloc(f"t{i}"))
    assert ("a",) not in c                               # un-pinned: evictable again
    tok2 = c.pin_all()
    tok3 = c.pin_all()                                   # pins nest (two graphs over the same module)
    c.unpin(tok2)
    assert all(c.pinned(k) for k in tok3)
    c.unpin(tok3)
    assert not any(c.pinned(k) for k in c.keys())
    c.clear()
    assert len(c) == 0


def _write_png_pairs(root, n, h=16, w=24, seed=0):
    from PIL import Image
    rng = np.random.default_rng(seed)
    hr_dir, lr_dir = os.path.join(root, "hr"), os.path.join(root, "lr")
    os.makedirs(hr_dir), os.makedirs(lr_dir)
    for i in range(n):
        name = f"sub-HC{i // 3:03d}_ses-01_T1map.nii_s{i:03d}.png"
        Image.fromarray(rng.integers(0, 256, (2 * h, 2 * w)).astype(np.uint8)).save(os.path.join(hr_dir, name))
        Image.fromarray(rng.integers(0, 256, (h, w)).astype(np.uint8)).save(os.path.join(lr_dir, name))
    return hr_dir, lr_dir


def test_png_pair_dataset_split_and_sharded_loader(tmp_path):
    """data.py against the reference's rules: pair discovery / subject ids (utils/dataset.py:37-67), the random_split partition
    for a seed (scripts/train.py:205-213), ragged last batch single-process, equal shards under data parallelism."""
    import torch
    from torch.utils.data import random_split
    from mri_superresolution_b200.data import PairedPNGDataset, UInt8BatchLoader, split_indices
    hr_dir, lr_dir = _write_png_pairs(str(tmp_path), 11)
    open(os.path.join(hr_dir, "notes.txt"), "w").close()
    ds = PairedPNGDataset(hr_dir, lr_dir)
    assert len(ds) == 11 and ds.subjects[0] == "HC000" and ds.subjects[10] == "HC003"
    assert ds.get_subject_indices("HC001") == [3, 4, 5]
    tr, va = split_indices(len(ds), 0.2, seed=42)
    ref_tr, ref_va = random_split(range(11), [9, 2], generator=torch.Generator().manual_seed(42))
    assert tr == list(ref_tr.indices) and va == list(ref_va.indices)
    low, full = ds.load(3)
    assert low.dtype == np.uint8 and low.shape == (16, 24) and full.shape == (32, 48)
    batches = list(UInt8BatchLoader(ds, tr, 4, shuffle=False, num_workers=2))
    assert [b[0].shape[0] for b in batches] == [4, 4, 1]                       # no drop_last, like the reference's DataLoader
    assert batches[0][0].dtype == torch.uint8 and batches[0][1].shape == (4, 32, 48)
    assert np.array_equal(batches[0][0][1].numpy(), ds.load(tr[1])[0])
    shards = [list(UInt8BatchLoader(ds, tr, 2, shuffle=True, seed=5, num_workers=0, rank=r, world=2)) for r in range(2)]
    assert len(shards[0]) == len(shards[1]) == 2                               # 9 pairs -> two global batches of 4, one pair dropped
    seen = torch.cat([b[0].reshape(b[0].shape[0], -1)[:, :8] for s in shards for b in s])
    assert len({tuple(r.tolist()) for r in seen}) == 8                         # disjoint shards
    e1 = [b[0] for b in UInt8BatchLoader(ds, tr, 4, shuffle=True, seed=1, num_workers=0)]
    e2 = [b[0] for b in UInt8BatchLoader(ds, tr, 4, shuffle=True, seed=1, num_workers=0)]
    assert all(torch.equal(a, b) for a, b in zip(e1, e2))                      # seeded order


def test_reference_import_names_resolve_to_the_drop_in_modules():
    """compat/: `from models.unet_model import UNetSuperRes`, `from utils.losses import CombinedLoss, SSIM` (scripts/train.py:26-33)
    give the B200 classes when the directory is first on sys.path."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys; sys.path.insert(0, %r); sys.path.insert(1, %r)\n"
            "from models.unet_model import UNetSuperRes\nfrom utils.losses import CombinedLoss, SSIM, ssim, PerceptualLoss\n"
            "from utils.dataset import MRISuperResDataset\n"
            "print(UNetSuperRes.__module__, CombinedLoss.__module__, SSIM.__module__)\n") % (
                os.path.join(root, "mri_superresolution_b200", "compat"), root)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd="/tmp")
    assert res.returncode == 0, res.stderr
    assert res.stdout.split() == ["mri_superresolution_b200.models.unet_model", "mri_superresolution_b200.utils.losses",
                                  "mri_superresolution_b200.utils.losses"]


def test_cli_argument_surface_matches_the_reference():
    """Same flags and defaults as scripts/train.py:486-548 and scripts/infer.py:452-490."""
    from mri_superresolution_b200 import infer as I
    from mri_superresolution_b200 import train as T
    a = T.parse_args(["--full_res_dir", "hr", "--low_res_dir", "lr"])
    assert (a.model_type, a.base_filters, a.batch_size, a.epochs, a.learning_rate, a.weight_decay, a.ssim_weight, a.perceptual_weight,
            a.vgg_layer_idx, a.perceptual_loss_type, a.initial_alpha, a.validation_split, a.patience, a.checkpoint_dir, a.log_dir) == (
        "unet", 32, 8, 100, 1e-4, 1e-5, 0.3, 0.0, 35, "l1", 0.0, 0.2, 10, "./checkpoints", "./logs")
    assert not (a.augmentation or a.use_tensorboard or a.use_amp or a.cpu) and 1 <= a.seed <= 10000
    b = I.parse_args(["--input", "x.png", "--output", "y.png"])
    assert (b.target, b.checkpoint_dir, b.checkpoint_path, b.model_type, b.base_filters) == (None, "./checkpoints", None, "unet", 64)
    with pytest.raises(SystemExit):
        T.parse_args(["--full_res_dir", "hr", "--low_res_dir", "lr", "--model_type", "cnn"])


def test_size_queries_survive_degenerate_arguments():
    """Every host-only size / layout query of include/mrisr.h with zero, negative, misaligned and oversized arguments: a number or
    0 comes back, the process does not die (an 8-pixel-wide slice used to divide by zero in the workspace arithmetic).  Run in a
    child process so that a crash fails this test instead of the session."""
    import subprocess
    import sys
    body = """
import ctypes as C, itertools
from mri_superresolution_b200 import _native
L = _native.lib()
for n, h, w, f, p, t in itertools.product((0, 1, -1, 3), (0, 8, 16, -8, 24, 2048), (0, 8, 16, 1024), (0, 16, 32, 48, 64, -16, 8), (0, 1, 2, 7), (0, 1)):
    L.mrisr_unet_workspace_bytes(n, h, w, f, p, t)
for f in (0, 8, 16, 32, 64, 128, -16, 24):
    L.mrisr_unet_param_elems(f)
    for i in (-1, 0, 5, 63, 64, 100):
        L.mrisr_unet_param_offset(f, i); L.mrisr_unet_param_size(f, i)
    for c in (-1, 0, 19, 20):
        L.mrisr_unet_grads_final_from(f, c)
for f, l, h, w in itertools.product((0, 16, 32, 64, -1), (-1, 0, 8, 19, 20), (0, 8, 256, 1024), (0, 8, 256, 1024)):
    L.mrisr_unet_layer_engine(f, l, h, w)
for n, h, w in itertools.product((0, 1, -1, 64), (0, 1, 31, 512, -3), (0, 1, 33, 512)):
    L.mrisr_l1_ssim_scratch_bytes(n, h, w); L.mrisr_image_metrics_scratch_bytes(n, h, w)
    L.mrisr_kspace_scratch_bytes(n, h, w); L.mrisr_robust_normalize_scratch_bytes(n, h, w)
    L.mrisr_match_histograms_scratch_bytes(n, h, w, h, w)
for n, h, w, c1, c2, co, ks in itertools.product((0, 1), (0, 1, 16), (0, 24), (0, 8, 16, 512, 24), (0, 16, 512), (0, 8, 256, 512, 24), (0, 1, 3, 5)):
    L.mrisr_conv2d_scratch_bytes(n, h, w, c1, c2, co, ks)
for cfg in ([], [64], [64, -1], [-1, 64], [64, 64, -1, 128], [7], [64] * 40, [64, -1, -1, 128], [0], [512, 512]):
    a = (C.c_int * max(1, len(cfg)))(*cfg)
    L.mrisr_vgg_weight_elems(a, len(cfg)); L.mrisr_vgg_packed_bytes(a, len(cfg))
    for n, h, w, g in itertools.product((0, 1, -1), (0, 16, 24, 256), (0, 16, 48), (0, 1)):
        L.mrisr_vgg_workspace_bytes(n, h, w, a, len(cfg), g); L.mrisr_vgg_workspace_bytes_fp32(n, h, w, a, len(cfg), g)
print("survived")
"""
    r = subprocess.run([sys.executable, "-c", body], capture_output=True, text=True, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0 and "survived" in r.stdout, (r.returncode, r.stderr[-500:])
