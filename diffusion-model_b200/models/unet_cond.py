"""Conditional U-Net noise predictor -- B200-native drop-in for the reference module
``models/unet_cond.py`` (class names, constructor arguments and state_dict keys are the
reference's: unet_cond.py:10-30, 32-52, 54-100, 102-217).

Forward is one call into libdmb200 (tcgen05 implicit-GEMM convolutions, fused GroupNorm /
GELU / embedding adds, attention); there is no PyTorch arithmetic and no CPU path.
"""
from __future__ import annotations

import torch
from torch import nn

from ._holders import Holder, double_conv, slots


class ResBlock(Holder):
    def __init__(self, in_channels, out_channels, mid_channels=None, residual=False):
        super().__init__()
        self.residual = residual
        self.double_conv = double_conv(in_channels, mid_channels or out_channels, out_channels)


class AttenionBlock(Holder):  # (sic) name kept for import compatibility
    def __init__(self, channels):
        super().__init__()
        self.channels = channels
        self.mha = nn.MultiheadAttention(channels, 4, batch_first=True)
        self.ln = nn.LayerNorm([channels])
        self.ff_self = slots({0: nn.LayerNorm([channels]), 1: nn.Linear(channels, channels),
                              3: nn.Linear(channels, channels)})


class Down(Holder):
    def __init__(self, in_channels, out_channels, emb_dim=256):
        super().__init__()
        self.maxpool_conv = slots({1: ResBlock(in_channels, in_channels, residual=True),
                                   2: ResBlock(in_channels, out_channels)})
        self.emb_layer = slots({1: nn.Linear(emb_dim, out_channels)})


class Up(Holder):
    def __init__(self, in_channels, out_channels, emb_dim=256):
        super().__init__()
        self.conv = slots({0: ResBlock(in_channels, in_channels, residual=True),
                           1: ResBlock(in_channels, out_channels, in_channels // 2)})
        self.emb_layer = slots({1: nn.Linear(emb_dim, out_channels)})


class UnetCond(nn.Module):
    """Constructor signature of the reference (unet_cond.py:113).  ``forward`` returns eps."""

    def __init__(self, in_ch=4, time_dim=256, num_classes=3, cfg_drop_prob=0.1, remove_deep_conv=False):
        super().__init__()
        self.in_ch = in_ch
        self.time_dim = time_dim
        self.remove_deep_conv = remove_deep_conv
        self.num_classes = num_classes
        self.cfg_drop_prob = cfg_drop_prob
        self.precision = None  # None -> DMB200_PRECISION env or "bf16"; else "bf16" / "tf32"

        self.class_emb = nn.Embedding(num_classes + 1, time_dim)
        self.cond_mlp = slots({0: nn.Linear(12 * 2, time_dim), 2: nn.Linear(time_dim, time_dim)})
        self.inc = ResBlock(in_ch, 64)
        self.down1 = Down(64, 128)
        self.sa1 = AttenionBlock(128)
        self.down2 = Down(128, 256)
        self.sa2 = AttenionBlock(256)
        self.down3 = Down(256, 256)
        self.sa3 = AttenionBlock(256)
        if remove_deep_conv:
            self.bot1 = ResBlock(256, 256)
            self.bot3 = ResBlock(256, 256)
        else:
            self.bot1 = ResBlock(256, 512)
            self.bot2 = ResBlock(512, 512)
            self.bot3 = ResBlock(512, 256)
        self.up1 = Up(512, 128)
        self.sa4 = AttenionBlock(128)
        self.up2 = Up(256, 64)
        self.sa5 = AttenionBlock(64)
        self.up3 = Up(128, 64)
        self.sa6 = AttenionBlock(64)
        self.out = nn.Conv2d(64, in_ch, kernel_size=1)

    # ------------------------------------------------------------------
    def _native(self, x: torch.Tensor):
        from dmb200 import context_for
        if self.in_ch != 4 or self.time_dim != 256 or self.num_classes != 3 or self.remove_deep_conv:
            raise NotImplementedError(
                "libdmb200 implements the configuration the sampling path uses: in_ch=4, time_dim=256, "
                "num_classes=3, remove_deep_conv=False")
        if x.dim() != 4 or tuple(x.shape[1:]) != (4, 28, 28):
            raise ValueError(f"latent must be (B,4,28,28), got {tuple(x.shape)}")
        dev = self.out.weight.device
        if x.device != dev:
            raise ValueError(f"input on {x.device} but parameters on {dev}")
        ctx = context_for(dev, self.precision)
        ctx.ensure_unet(self)
        return ctx

    # ---- weight-change tracking for the native copy (dmb200.native.Context._params)
    _native_epoch = 0

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._native_epoch += 1
        return out

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self._native_epoch += 1
        return out

    def refresh_native(self):
        """Rebuild the repacked device copy of the weights at the next call.  load_state_dict,
        Utils.loadModel and .to() are detected automatically; call this after writing parameters
        through `.data` (no version counter moves)."""
        from dmb200.native import invalidate_module
        self._native_epoch += 1
        invalidate_module(self)

    def _run(self, x, t, y, cond_vals, cond_mask, want_geom):
        ctx = self._native(x)
        if (cond_vals is None) or (cond_mask is None):
            cond_vals = cond_mask = None  # the reference adds the numeric term only when both are given
        if bool(((y < 0) | (y > self.num_classes)).any()):
            raise IndexError("index out of range in self")     # what nn.Embedding raises (unet_cond.py:121)
        with torch.no_grad():
            return ctx.unet_forward(x.float(), t, y, cond_vals, cond_mask, want_geom=want_geom)

    def forward(self, x, t, y, cond_vals=None, cond_mask=None, cond_drop_prob=None):
        if self.training and self.cfg_drop_prob > 0:
            raise NotImplementedError("training-mode label dropping (unet_cond.py:199-211) is outside the "
                                      "sampling path; call .eval()")
        eps, _ = self._run(x, t, y, cond_vals, cond_mask, want_geom=False)
        return eps
