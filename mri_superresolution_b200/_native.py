"""ctypes binding of libmrisr_b200.so (the C ABI declared in include/mrisr.h).

There is no CPU fallback and no alternative backend: if the library is missing or a tensor is
not on a CUDA device the callers raise.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmrisr_b200.so")
CSRC = os.path.join(_HERE, "csrc")

FP32 = 0
BF16 = 1

_lib = None

# name -> (restype, argtypes); must list every symbol include/mrisr.h declares
SIGNATURES = {
    "mrisr_version": (c_int, []),
    "mrisr_last_error": (c_char_p, []),
    "mrisr_unet_param_tensors": (c_int, [c_int]),
    "mrisr_unet_param_elems": (c_int64, [c_int]),
    "mrisr_unet_param_offset": (c_int64, [c_int, c_int]),
    "mrisr_unet_param_size": (c_int64, [c_int, c_int]),
    "mrisr_unet_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "mrisr_unet_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_unet_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                    c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_unet_grads_final_from": (c_int64, [c_int, c_int]),
    "mrisr_unet_backward_marked": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                           c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "mrisr_event_create": (c_void_p, []),
    "mrisr_event_destroy": (c_int, [c_void_p]),
    "mrisr_stream_wait_event": (c_int, [c_void_p, c_void_p]),
    "mrisr_l1_ssim_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mrisr_l1_ssim_fwd_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                      c_float, c_float, c_int, c_float, c_float, c_void_p]),
    "mrisr_ssim_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                   c_int, c_float, c_float, c_void_p]),
    "mrisr_ssim_static_scratch_bytes": (c_size_t, []),
    "mrisr_ssim_forward_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_float, c_void_p]),
    "mrisr_u8_to_float": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "mrisr_percentile_normalize_u8": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float, c_void_p]),
    "mrisr_augment_u8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "mrisr_image_metrics_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mrisr_image_metrics": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "mrisr_upscale2x_u8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_match_histograms_scratch_bytes": (c_size_t, [c_int] * 5),
    "mrisr_match_histograms": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_kspace_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mrisr_simulate_low_field": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int,
                                         c_double, c_double, c_uint64, c_void_p]),
    "mrisr_area_half_f64": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_robust_normalize_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mrisr_robust_normalize_f64": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int,
                                           c_double, c_double, c_double, c_double, c_void_p]),
    "mrisr_unet_debug_tensor": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                        c_void_p, c_void_p]),
    "mrisr_planar_geometry": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p]),
    "mrisr_conv2d_scratch_bytes": (c_size_t, [c_int] * 7),
    "mrisr_conv2d_bf16": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_size_t, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_conv2d_wgrad_bf16": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_size_t,
                                        c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mrisr_vgg_weight_elems": (c_int64, [c_void_p, c_int]),
    "mrisr_vgg_packed_bytes": (c_size_t, [c_void_p, c_int]),
    "mrisr_vgg_pack_weights": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "mrisr_vgg_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_void_p, c_int, c_int]),
    "mrisr_vgg_features": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int,
                                   c_void_p, c_int, c_void_p]),
    "mrisr_vgg_perceptual_fwd_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                             c_int, c_int, c_int, c_void_p, c_int, c_int, c_float, c_void_p]),
    "mrisr_vgg_workspace_bytes_fp32": (c_size_t, [c_int, c_int, c_int, c_void_p, c_int, c_int]),
    "mrisr_vgg_features_fp32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int,
                                        c_void_p, c_int, c_void_p]),
    "mrisr_vgg_perceptual_fwd_bwd_fp32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                                  c_int, c_int, c_int, c_void_p, c_int, c_int, c_float, c_void_p]),
    "mrisr_unet_layer_engine": (c_int, [c_int, c_int, c_int, c_int]),
    "mrisr_launch_count": (ctypes.c_longlong, []),
    "mrisr_prof_enable": (c_int, [c_int]),
    "mrisr_prof_report": (c_int, [c_char_p, c_size_t]),
    "mrisr_adam_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
                                c_float, c_float, c_int, c_float, c_void_p]),
    "mrisr_adam_step_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
                                    c_float, c_void_p, c_float, c_void_p]),
}


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into lib/libmrisr_b200.so (nvcc cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", CSRC, "-j", str(os.cpu_count() or 4)], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libmrisr_b200.so failed")
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU or PyTorch fallback for this path)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc == 0:
        return
    msg = lib().mrisr_last_error().decode("utf-8", "replace")
    if rc < 0:
        raise ValueError(f"{what}: {msg} (code {rc})")
    raise RuntimeError(f"{what}: CUDA error {rc}: {msg}")


def ptr(t) -> int:
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("mrisr kernels need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("mrisr kernels need contiguous tensors")
    return t.data_ptr()


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def launch_count() -> int:
    """Kernel launches issued by the library in this process so far."""
    return int(lib().mrisr_launch_count())


def profile(on: bool) -> None:
    _native_check = lib().mrisr_prof_enable(1 if on else 0)
    assert _native_check == 0


def profile_report():
    """[(tag, calls, total_ms)] since profiling was enabled (synchronises the device)."""
    buf = ctypes.create_string_buffer(1 << 16)
    lib().mrisr_prof_report(buf, len(buf))
    out = []
    for line in buf.value.decode().splitlines():
        tag, calls, ms = line.rsplit(" ", 2)
        out.append((tag, int(calls), float(ms)))
    return out
