"""Import the UNMODIFIED reference from /root/reference (build container only).

TEST INFRASTRUCTURE ONLY.  ``/root/reference`` does not exist on the GPU box, so
nothing in the ``-m gpu`` tests, ``smoke()`` or ``bench.py`` calls this at run
time; it is used by ``tests/golden/make_golden.py`` and by the CPU tests that
cross-check the restatement against the live reference (skipped when absent).

The only accommodation is a ``matplotlib`` stub: diff.py:2-3 and utils.py:4
import matplotlib (never called on the sampling path) and it is not installed
here.
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("DMB_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "diff.py"))


def _stub_matplotlib() -> None:
    if "matplotlib" in sys.modules:
        return
    try:
        import matplotlib  # noqa: F401
        return
    except Exception:
        pass
    m = types.ModuleType("matplotlib")
    m.use = lambda *a, **k: None
    sys.modules["matplotlib"] = m
    sys.modules["matplotlib.pyplot"] = types.ModuleType("matplotlib.pyplot")


_PRODUCT_NAMES = ("diff", "entityCsvSampler", "utils", "models", "models.unet_cond",
                  "models.unet_cond_geom", "models.vae", "models.unet", "eval_iou_noise")


@contextlib.contextmanager
def reference_modules():
    """Context manager yielding a namespace with the reference's own modules.

    The reference and the product use the same top-level module names (that is the
    drop-in boundary), so the product's entries are moved out of ``sys.modules`` for
    the duration and restored afterwards.
    """
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    _stub_matplotlib()
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k in _PRODUCT_NAMES}
    saved_path = list(sys.path)
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import diff, entityCsvSampler, eval_iou_noise  # noqa: E401
        from models import unet_cond, unet_cond_geom, vae
        ns = types.SimpleNamespace(diff=diff, entityCsvSampler=entityCsvSampler,
                                   eval_iou_noise=eval_iou_noise, unet_cond=unet_cond,
                                   unet_cond_geom=unet_cond_geom, vae=vae)
        yield ns
    finally:
        for k in list(sys.modules):
            if k in _PRODUCT_NAMES:
                del sys.modules[k]
        sys.modules.update(saved)
        sys.path[:] = saved_path


@contextlib.contextmanager
def injected_noise(x_T, noise_seq):
    """Replace torch.randn / torch.randn_like while the reference samples, so its
    trajectory uses pre-generated tensors (diff.py:158, 327 draw from the global
    generator; the reference has no seed handling).  ``noise_seq`` is an iterator
    of per-step noise tensors in call order."""
    import torch
    real_randn, real_like = torch.randn, torch.randn_like
    it = iter(noise_seq)

    def fake_randn(*a, **k):
        return x_T.clone()

    def fake_like(x, **k):
        return next(it).clone()

    torch.randn, torch.randn_like = fake_randn, fake_like
    try:
        yield
    finally:
        torch.randn, torch.randn_like = real_randn, real_like
