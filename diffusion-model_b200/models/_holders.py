"""Parameter holders.

The modules in this package own fp32 parameters in the reference's state_dict layout
(so ``Utils.loadModel`` / ``load_state_dict`` work on reference checkpoints) but compute
nothing themselves: the arithmetic runs in libdmb200 on a repacked device copy.  Leaf
``torch.nn`` layers are used purely as containers, created in the reference's construction
order so that ``torch.manual_seed(s); Model()`` yields the same random-init weights.
"""
from __future__ import annotations

from typing import Dict

from torch import nn


def slots(entries: Dict[int, nn.Module]) -> nn.ModuleDict:
    """Children addressed by integer position (``<name>.<i>.weight``), skipping the
    parameter-free positions an ``nn.Sequential`` of the reference would have."""
    return nn.ModuleDict({str(i): m for i, m in entries.items()})


def double_conv(c_in: int, c_mid: int, c_out: int) -> nn.ModuleDict:
    """conv3x3 / GroupNorm(1) / [GELU] / conv3x3 / GroupNorm(1)  (models/unet_cond.py:18-24)."""
    return slots({
        0: nn.Conv2d(c_in, c_mid, kernel_size=3, padding=1, bias=False),
        1: nn.GroupNorm(1, c_mid),
        3: nn.Conv2d(c_mid, c_out, kernel_size=3, padding=1, bias=False),
        4: nn.GroupNorm(1, c_out),
    })


class Holder(nn.Module):
    """A module that only groups parameters; calling it is an error by design."""

    def forward(self, *a, **k):  # pragma: no cover - guard
        raise RuntimeError(
            f"{type(self).__name__} holds parameters only; the network runs as a whole in "
            "libdmb200 (call the top-level model)")
