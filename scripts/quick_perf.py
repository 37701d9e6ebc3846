"""Throw-away timing probe: ms per CFG step at several batch sizes, decode time."""
import sys, os, time
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
prec = os.environ.get("DMB200_PRECISION", "bf16")
ctx = context_for(dev, prec); ctx.ensure_unet(unet); ctx.ensure_vae(vae)
d = Diffuser(device=dev); coef = d.step_coefficients()
for B in [int(a) for a in sys.argv[1:]] or [1, 16, 256]:
    y, vals, mask, _ = synth.mixed_conditions(B, 0)
    y, vals, mask = y.to(dev), vals.to(dev), mask.to(dev)
    x = torch.randn((B, 4, 28, 28), device=dev)
    n = 20
    ctx.sample_loop(x, y, 0, vals, mask, 3.0, coef, 1000, n, seed=1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); ctx.sample_loop(x, y, 0, vals, mask, 3.0, coef, 980, n, seed=1); e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    z = torch.randn((B, 4, 28, 28), device=dev)
    ctx.vae_decode(z, want_f32=False, want_u8=True); torch.cuda.synchronize()
    e0.record(); ctx.vae_decode(z, want_f32=False, want_u8=True); e1.record(); torch.cuda.synchronize()
    dms = e0.elapsed_time(e1)
    tot = ms * 1000 + dms
    print(f"{prec} B={B}: {ms:.3f} ms/step -> T=1000 {ms:.1f}s + decode {dms:.1f} ms -> {B / tot * 1000:.2f} samples/s "
          f"({B / tot * 1000 * 5.4066:.1f} TFLOP/s algorithmic)", flush=True)
