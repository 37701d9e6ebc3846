"""VAE -- B200-native drop-in for ``models/vae.py`` (vae.py:5-75).

``decode`` is on the sampling path (vae.py:64-69); it runs in libdmb200 (tcgen05
convolutions / transposed convolutions, GroupNorm(8)+GELU, sigmoid).  ``encode``
(vae.py:51-62) runs there too: the strided 4x4 convolutions become
2x2-tap GEMMs over a space-to-depth activation, and the 1x1 heads, the clamp, the
reparameterisation and the KL term are one kernel.  ``forward`` (vae.py:71-76) = encode, decode and
the reconstruction error (dmb_recon_error).  Forward only: no autograd."""
from __future__ import annotations

import torch
from torch import nn

from ._holders import slots


class VAE(nn.Module):
    def __init__(self, in_channels=3, z_channels=4, base_channels=64, scale_factor=0.18215):
        super().__init__()
        self.in_channels = in_channels
        self.z_channels = z_channels
        self.base_channels = base_channels
        self.scale_factor = scale_factor
        self.precision = None
        b = base_channels
        self.enc = slots({
            0: nn.Conv2d(in_channels, b, 3, stride=1, padding=1), 1: nn.GroupNorm(8, b),
            3: nn.Conv2d(b, b, 4, stride=2, padding=1), 4: nn.GroupNorm(8, b),
            6: nn.Conv2d(b, b * 2, 3, stride=1, padding=1), 7: nn.GroupNorm(8, b * 2),
            9: nn.Conv2d(b * 2, b * 2, 4, stride=2, padding=1), 10: nn.GroupNorm(8, b * 2),
            12: nn.Conv2d(b * 2, b * 4, 3, stride=1, padding=1), 13: nn.GroupNorm(8, b * 4),
            15: nn.Conv2d(b * 4, b * 4, 4, stride=2, padding=1), 16: nn.GroupNorm(8, b * 4),
        })
        self.to_mu = nn.Conv2d(b * 4, z_channels, 1)
        self.to_logvar = nn.Conv2d(b * 4, z_channels, 1)
        self.dec = slots({
            0: nn.Conv2d(z_channels, b * 4, 3, padding=1), 1: nn.GroupNorm(8, b * 4),
            3: nn.ConvTranspose2d(b * 4, b * 4, 4, stride=2, padding=1), 4: nn.GroupNorm(8, b * 4),
            6: nn.Conv2d(b * 4, b * 2, 3, padding=1), 7: nn.GroupNorm(8, b * 2),
            9: nn.ConvTranspose2d(b * 2, b * 2, 4, stride=2, padding=1), 10: nn.GroupNorm(8, b * 2),
            12: nn.Conv2d(b * 2, b, 3, padding=1), 13: nn.GroupNorm(8, b),
            15: nn.ConvTranspose2d(b, b, 4, stride=2, padding=1), 16: nn.GroupNorm(8, b),
            18: nn.Conv2d(b, in_channels, 3, padding=1),
        })

    def _ctx(self, z: torch.Tensor):
        from dmb200 import context_for
        if (self.in_channels, self.z_channels, self.base_channels) != (3, 4, 64) or abs(self.scale_factor - 0.18215) > 1e-12:
            raise NotImplementedError("libdmb200 implements the default VAE (3, 4, 64, scale 0.18215)")
        if z.dim() != 4 or tuple(z.shape[1:]) != (4, 28, 28):
            raise ValueError(f"latent must be (B,4,28,28), got {tuple(z.shape)}")
        dev = self.dec["18"].weight.device
        if z.device != dev:
            raise ValueError(f"input on {z.device} but parameters on {dev}")
        ctx = context_for(dev, self.precision)
        ctx.ensure_vae(self)
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
        """Rebuild the device copy of the weights at the next call (see UnetCond.refresh_native)."""
        from dmb200.native import invalidate_module
        self._native_epoch += 1
        invalidate_module(self)

    def decode(self, z):
        """z / scale -> decoder -> sigmoid; (B,3,224,224) fp32 in [0,1]  (vae.py:64-69)."""
        with torch.no_grad():
            img, _ = self._ctx(z).vae_decode(z, want_f32=True, want_u8=False)
        return img

    def decode_u8(self, z):
        """decode + the x*255 / clamp / truncating uint8 cast of Diffuser.reverse_to_img
        (diff.py:58-61), done in the decoder's epilogue."""
        with torch.no_grad():
            _, u8 = self._ctx(z).vae_decode(z, want_f32=False, want_u8=True)
        return u8

    def _enc_ctx(self, x: torch.Tensor):
        from dmb200 import context_for
        if (self.in_channels, self.z_channels, self.base_channels) != (3, 4, 64) or abs(self.scale_factor - 0.18215) > 1e-12:
            raise NotImplementedError("libdmb200 implements the default VAE (3, 4, 64, scale 0.18215)")
        if x.dim() != 4 or tuple(x.shape[1:]) != (3, 224, 224):
            raise ValueError(f"image batch must be (B,3,224,224), got {tuple(x.shape)}")
        dev = self.to_mu.weight.device
        if x.device != dev:
            raise ValueError(f"input on {x.device} but parameters on {dev}")
        ctx = context_for(dev, self.precision)
        ctx.ensure_vae_enc(self)
        return ctx

    def encode(self, x, eps=None):
        """(z, kl.mean()) like the reference (vae.py:51-62).  ``eps`` defaults to
        ``torch.randn(B,4,28,28)`` drawn from torch's generator on the module's device -- the same
        draw ``torch.randn_like(std)`` makes in the reference (vae.py:56)."""
        with torch.no_grad():
            ctx = self._enc_ctx(x)
            if eps is None:
                eps = torch.randn((x.shape[0], 4, 28, 28), device=x.device, dtype=torch.float32)
            z, kl, _, _ = ctx.vae_encode(x, eps)
            return z, kl.mean()

    def encode_moments(self, x, eps=None):
        """encode() plus the posterior moments: (z, kl_per_sample, mu, logvar_clamped)."""
        with torch.no_grad():
            ctx = self._enc_ctx(x)
            if eps is None:
                eps = torch.randn((x.shape[0], 4, 28, 28), device=x.device, dtype=torch.float32)
            return ctx.vae_encode(x, eps, want_moments=True)

    def forward(self, x, eps=None):
        """encode -> decode -> loss, forward only (vae.py:71-76): returns
        (x_recon, z, loss, {'recon_mse', 'kl'}) with loss = mse(x_recon, x) + 1e-6 * kl.  What
        vae_recon_check.py and a validation pass need; there are no backward kernels, so a call that
        would have to record a graph (grad mode on and a parameter that requires grad, i.e. a
        training step) raises instead of returning a loss that cannot be differentiated."""
        if torch.is_grad_enabled() and self.training and any(p.requires_grad for p in self.parameters()):
            raise NotImplementedError("VAE.forward in training mode needs backward kernels, which libdmb200 does not "
                                      "have (forward only): call it under torch.no_grad() or after .eval()")
        with torch.no_grad():
            z, kl = self.encode(x, eps)
            x_recon = self.decode(z)
            sq, _ = self._enc_ctx(x).recon_error(x_recon, x.to(torch.float32))
            recon = (sq.sum() / float(x.numel())).to(torch.float32)
            loss = recon + 1e-6 * kl
        return x_recon, z, loss, {"recon_mse": recon, "kl": kl}
