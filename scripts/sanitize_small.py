"""A small pass over the round-2 kernels for compute-sanitizer (memcheck / racecheck / initcheck):
one CFG step at batch 2 and 5, a decode of 3 and of 33 images (the narrowed tiles and the full chunk
plus a tail plan), an encode -> decode -> reconstruction error, the scorer, in bf16 and tf32 storage.
usage: [compute-sanitizer --tool memcheck] python scripts/sanitize_small.py
(compute-sanitizer is closed on the GPU pool this was developed on; the plain run is what was exercised)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "diffusion-model_b200")]
import torch
from oracle import synth
from models.unet_cond_geom import UnetCondWithGeomHead
from models.vae import VAE
from diff import Diffuser
from dmb200 import context_for

dev = torch.device("cuda", 0)
torch.manual_seed(0)
unet = UnetCondWithGeomHead().eval().to(dev); vae = VAE().eval().to(dev)
coef = Diffuser(device=dev).step_coefficients()
for prec in ("bf16", "tf32"):
    ctx = context_for(dev, prec); ctx.ensure_unet(unet); ctx.ensure_vae(vae); ctx.ensure_vae_enc(vae)
    for B in (2, 5):
        y, vals, mask, _ = synth.mixed_conditions(B, 0)
        y, vals, mask = y.to(dev), vals.to(dev), mask.to(dev)
        x = torch.randn((B, 4, 28, 28), device=dev)
        t = torch.full((B,), 500, dtype=torch.long, device=dev)
        x = ctx.denoise_cond_step(x, t, y, 0, vals, mask, 3.0, coef, seed=1)
        ctx.sample_loop(x, y, 0, vals, mask, 3.0, coef, 20, 12, seed=2)
    for n in (3, 33):
        z = torch.randn((n, 4, 28, 28), device=dev)
        img, u8 = ctx.vae_decode(z, want_f32=True, want_u8=True)
    xin = torch.rand((3, 3, 224, 224), device=dev)
    zz, kl, _, _ = ctx.vae_encode(xin, torch.randn((3, 4, 28, 28), device=dev))
    rec, _ = ctx.vae_decode(zz)
    sq, ab = ctx.recon_error(rec, xin)
    cnt, hit = ctx.score_masks(u8[:3], u8[:3].flip(0), threshold=128, invert=True, sigma=2.0)
    torch.cuda.synchronize()
    print(prec, "ok", float(x.abs().max()), float(sq.sum()), int(cnt.sum()))
