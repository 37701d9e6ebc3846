"""UnetCondWithGeomHead -- B200-native drop-in for ``models/unet_cond_geom.py``
(unet_cond_geom.py:8-23, 26-100).  ``forward`` returns ``(eps_pred, geom_pred)``."""
from __future__ import annotations

import torch
from torch import nn

from ._holders import Holder, slots
from .unet_cond import AttenionBlock, Down, ResBlock, UnetCond, Up  # noqa: F401  (re-exported like the reference)


class GeomHead(Holder):
    def __init__(self, in_ch: int, out_dim: int, hidden: int = 256):
        super().__init__()
        self.mlp = slots({0: nn.Linear(in_ch, hidden), 2: nn.Linear(hidden, out_dim)})


class UnetCondWithGeomHead(UnetCond):
    def __init__(self, in_ch=4, time_dim=256, num_classes=3, cfg_drop_prob=0.0, remove_deep_conv=False,
                 geom_dim=12, geom_hidden=256):
        super().__init__(in_ch=in_ch, time_dim=time_dim, num_classes=num_classes, cfg_drop_prob=cfg_drop_prob,
                         remove_deep_conv=remove_deep_conv)
        self.geom_dim = geom_dim
        self.geom_hidden = geom_hidden
        self.geom_head = GeomHead(in_ch=64, out_dim=geom_dim, hidden=geom_hidden)

    def forward(self, x: torch.Tensor, t: torch.Tensor, y: torch.Tensor, cond_vals: torch.Tensor = None,
                cond_mask: torch.Tensor = None, cond_drop_prob: float = 0.0):
        if self.geom_dim != 12 or self.geom_hidden != 256:
            raise NotImplementedError("libdmb200 implements geom_dim=12, geom_hidden=256")
        return self._run(x, t, y, cond_vals, cond_mask, want_geom=True)
