"""mrisr_ssim_forward_static: the allocation-free SSIM metric entry point (include/mrisr.h; no counterpart in the reference).
Called through the C ABI and compared with mrisr_ssim_forward (caller-owned
scratch) on two calls in a row with different shapes and data -- the static array keeps the previous call's contents, which
must not leak into the next result.  Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(a, b):
    from mri_superresolution_b200 import _native
    L = _native.lib()
    n, _, h, w = a.shape
    st = _native.stream_ptr()
    scratch = torch.empty(L.mrisr_l1_ssim_scratch_bytes(n, h, w), dtype=torch.uint8, device=a.device)
    ref = torch.empty(n, dtype=torch.float32, device=a.device)
    _native.check(L.mrisr_ssim_forward(_native.ptr(a), _native.ptr(b), _native.ptr(ref), _native.ptr(scratch), n, h, w,
                                       11, 1.5, 1.0, st), "mrisr_ssim_forward")
    out = torch.full((n,), float("nan"), dtype=torch.float32, device=a.device)
    before = torch.cuda.memory_allocated(a.device)
    _native.check(L.mrisr_ssim_forward_static(_native.ptr(a), _native.ptr(b), _native.ptr(out), n, h, w, 11, 1.5, 1.0, st),
                  "mrisr_ssim_forward_static")
    assert torch.cuda.memory_allocated(a.device) == before          # nothing allocated through torch either
    torch.cuda.synchronize()
    return out, ref


def test_static_scratch_ssim_matches_the_scratch_entry_point_twice_in_a_row():
    from mri_superresolution_b200 import _native
    from oracle.synth import make_pairs
    L = _native.lib()
    dev = torch.device("cuda:0")
    assert L.mrisr_ssim_static_scratch_bytes() == 1 << 20
    g = torch.Generator("cpu").manual_seed(11)
    # call 1: a larger problem (fills more of the array); call 2: a smaller, ragged one on other data
    _, hr = make_pairs(6, 96, 80, seed=3)                           # 6 x 1 x 192 x 160
    a1 = (hr + 0.05 * torch.randn(hr.shape, generator=g)).clamp(0, 1).to(dev)
    o1, r1 = _both(a1, hr.to(dev))
    assert torch.allclose(o1, r1, rtol=0, atol=1e-6), (o1, r1)      # same kernels, same order: equal (test tolerance 1e-5)
    _, hr2 = make_pairs(3, 20, 28, seed=4)                          # 3 x 1 x 40 x 56
    a2 = (hr2 * 0.8 + 0.1).to(dev)
    o2, r2 = _both(a2, hr2.to(dev))
    assert torch.allclose(o2, r2, rtol=0, atol=1e-6), (o2, r2)
    assert float((o1[:3] - o2).abs().min()) > 1e-3                  # the second result is not the first call's leftovers
    # against the public metric, too (utils/losses.ssim = per-sample means of the same kernel)
    from mri_superresolution_b200.utils.losses import ssim
    assert torch.allclose(o2, ssim(a2, hr2.to(dev), size_average=False), atol=1e-5)


def test_static_scratch_ssim_rejects_sizes_beyond_the_array():
    from mri_superresolution_b200 import _native
    L = _native.lib()
    dev = torch.device("cuda:0")
    x = torch.zeros(1, 1, 32, 32, device=dev)
    out = torch.zeros(1, device=dev)
    # 4096 x 4096 slices: 16 384 tiles each; 9 of them need 1.125 MiB > 1 MiB (pointers are not dereferenced on rejection)
    rc = L.mrisr_ssim_forward_static(_native.ptr(x), _native.ptr(x), _native.ptr(out), 9, 4096, 4096, 11, 1.5, 1.0,
                                     _native.stream_ptr())
    assert rc == -2
    msg = L.mrisr_last_error().decode()
    assert "static scratch" in msg and "1048576" in msg, msg
    with pytest.raises(ValueError):
        _native.check(rc, "mrisr_ssim_forward_static")
