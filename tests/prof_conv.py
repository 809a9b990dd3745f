"""Runs one convolution building block a few times (profiling aid for ncu / CUDA-event timing on the GPU box):
PYTHONPATH=. python tests/prof_conv.py <engine 0|1|2> N H W C1 Cout [fwd|wgrad] [reps]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from mri_superresolution_b200 import _native


def main():
    eng, n, h, w, c1, cout = [int(v) for v in sys.argv[1:7]]
    what = sys.argv[7] if len(sys.argv) > 7 else "fwd"
    reps = int(sys.argv[8]) if len(sys.argv) > 8 else 5
    dev = torch.device("cuda:0")
    L = _native.lib()
    rows = n * (h + 2) * (w + 2)
    if eng == 2:
        g, pl = ctypes.c_longlong(), ctypes.c_longlong()
        L.mrisr_planar_geometry(n, h, w, ctypes.addressof(g), ctypes.addressof(pl))
        mk = lambda c: (torch.randn(c // 8, pl.value, 8, device=dev) * 0.5).to(torch.bfloat16)
        mkz = lambda c: torch.zeros(c // 8, pl.value, 8, device=dev, dtype=torch.bfloat16)
    else:
        mk = lambda c: (torch.randn(rows, c, device=dev) * 0.5).to(torch.bfloat16)
        mkz = lambda c: torch.zeros(rows, c, device=dev, dtype=torch.bfloat16)
    x, dst = mk(c1), mkz(cout)
    wt = torch.randn(cout, c1, 3, 3, device=dev) * 0.05
    dw = torch.zeros_like(wt)
    stats = torch.zeros(n, 8, 2, device=dev)
    sb = L.mrisr_conv2d_scratch_bytes(n, h, w, c1, 0, cout, 3)
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    for i in range(reps):
        ev[i].record()
        if what == "fwd":
            rc = L.mrisr_conv2d_bf16(_native.ptr(x), c1, None, 0, _native.ptr(wt), None, _native.ptr(dst), _native.ptr(stats),
                                     _native.ptr(scratch), sb, n, h, w, cout, 3, 0, eng, _native.stream_ptr())
        else:
            rc = L.mrisr_conv2d_wgrad_bf16(_native.ptr(x), c1, None, 0, _native.ptr(dst), _native.ptr(dw), _native.ptr(scratch), sb,
                                           n, h, w, cout, 3, eng, _native.stream_ptr())
        _native.check(rc, what)
    ev[reps].record()
    torch.cuda.synchronize()
    ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(reps)]
    fl = 2.0 * c1 * cout * 9 * h * w * n
    print(f"engine {eng} {what} N={n} {h}x{w} {c1}>{cout}: ms per call (incl. pack/finalize) {['%.3f' % t for t in ts]}  best {fl / min(ts) / 1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    main()
