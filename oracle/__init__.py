"""CPU oracle for the U-Net super-resolution hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it, and there only as the checker / the timed CPU
baseline -- never as the thing shipped.  The product path
(``mri_superresolution_b200``) never imports this package and raises when its CUDA
library is missing.

The oracle is a *functional* plain-PyTorch fp32 restatement of the reference's
``models/unet_model.py`` and ``utils/losses.py`` (every function cites the reference
file:line it follows).  Parity pinning: the reference ships no golden vectors or tests
of its own, so the oracle is pinned against outputs of the
reference modules themselves, imported from /root/reference in the build container by
``oracle/make_golden.py`` and committed under ``tests/golden/``.
"""
