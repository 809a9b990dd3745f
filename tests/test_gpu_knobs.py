"""A/B consistency of the kernel-selection knobs at a mid-size shape (f = 32, 128x128 -> 256x256, batch 3): every
alternative kernel for the same layer (fused / unfused GroupNorm backward, merged-tap / per-tap tensor-core convolution
and weight gradient, one / two CTAs per SM in the planar kernels, row-structured / ring weight gradient) must produce
the same training-step gradients up to bf16 accumulation order.  The golden-vector parity tests run at 32x32, where
several of these kernels degenerate (single cluster rank, one tile); this is the test that exercises them at a size
where the clusters, strips and two-CTA plans are really in use.  Each variant runs in a fresh process because the
library reads its MRISR_* knobs once."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHAPE = ("32", "3", "128", "128")


def _probe(tmp_path, name, env_extra):
    path = os.path.join(str(tmp_path), name + ".npz")
    env = dict(os.environ)
    env.update(env_extra)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "_grad_probe.py"), path, *SHAPE], env=env, cwd=ROOT,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    return dict(np.load(path))


@pytest.fixture(scope="module")
def baseline(tmp_path_factory):
    return _probe(tmp_path_factory.mktemp("knobs"), "default", {})


# (name, environment, forward_changes): knobs that only swap a BACKWARD kernel leave the forward pass bit-identical, so the
# gradients may differ by fp32 accumulation order only; knobs that swap a FORWARD kernel re-round some bf16 activations
# (max-pool ties, LeakyReLU signs flip for a few elements), which the cancellation-heavy GroupNorm parameter gradients amplify.
KNOBS = [
    ("gn_bwd_unfused", {"MRISR_GNBWD_FUSED": "0"}, False),
    ("wgrad_tc_per_tap", {"MRISR_WGRAD_TC3": "0"}, False),
    ("rows_one_cta", {"MRISR_ROWS_CTAS": "1"}, False),
    ("upbwd_tiled", {"MRISR_UPBWD_RING": "0"}, False),
    ("ssim_stream", {"MRISR_SSIM_STREAM": "1"}, False),                            # the streamed one-kernel loss                            # the tiled bilinear adjoint instead of the row ring
    # weight gradients on the main stream instead of the side stream: same kernels, same inputs (bit-equal, checked below)
    ("bwd_one_stream", {"MRISR_BWD_SIDE": "0"}, False),
    # the fused halo-2R loss kernel instead of the forward-maps + gradient pair
    ("ssim_fused", {"MRISR_SSIM_SPLIT": "0"}, False),
    ("conv_tc_per_tap", {"MRISR_CONV_TC_MERGE3": "0"}, True),
    ("ring_one_cta", {"MRISR_RING_CTAS": "1"}, True),
    # the wide-image fallback (unet_plan.cuh: level_is_planar): levels <= 0 in NHWC on the TMA kernels, as at 512 x 512 input
    ("levels_nhwc", {"MRISR_FORCE_NHWC": "1"}, True),
]


@pytest.mark.parametrize("name,env,fwd", KNOBS, ids=[k[0] for k in KNOBS])
def test_knob_variant_matches_default(tmp_path, baseline, name, env, fwd):
    v = _probe(tmp_path, name, env)
    assert abs(float(v["loss"]) - float(baseline["loss"])) <= (2e-3 if fwd else 1e-6) * abs(float(baseline["loss"]))
    worst = {}
    for k, b in baseline.items():
        if k == "loss":
            continue
        a = v[k].astype(np.float64).ravel()
        b = b.astype(np.float64).ravel()
        rel = np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-12)
        kind = "vector" if v[k].ndim <= 1 else "tensor"
        if rel >= worst.get(kind, (0.0, ""))[0]:
            worst[kind] = (rel, k)
        # measured on B200: backward-only knobs <= 2.2e-3 (dr is re-rounded to bf16 after a differently ordered reduction),
        # forward knobs <= 2.9e-2 on weights and 3.2e-2 on GroupNorm / bias vectors
        tol = (0.2 if kind == "vector" else 6e-2) if fwd else 5e-3
        assert rel < tol, (name, k, rel)
        if name == "bwd_one_stream":
            assert np.array_equal(v[k], baseline[k]), k          # only the stream changes: identical to the bit
    print(name, {k: (round(r, 6), n) for k, (r, n) in worst.items()})
