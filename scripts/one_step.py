"""One eager CFG step at batch B (for ncu captures).  usage: python scripts/one_step.py [B] [n_steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "diffusion-model_b200")]
import torch
from oracle import synth
from models.unet_cond_geom import UnetCondWithGeomHead
from diff import Diffuser
from dmb200 import context_for
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda", 0)
torch.manual_seed(0)
unet = UnetCondWithGeomHead().eval().to(dev)
ctx = context_for(dev, "bf16"); ctx.ensure_unet(unet)
coef = Diffuser(device=dev).step_coefficients()
y, vals, mask, _ = synth.mixed_conditions(B, 0)
y, vals, mask = y.to(dev), vals.to(dev), mask.to(dev)
x = torch.randn((B, 4, 28, 28), device=dev)
t = torch.full((B,), 500, dtype=torch.long, device=dev)
for _ in range(n):
    x = ctx.denoise_cond_step(x, t, y, 0, vals, mask, 3.0, coef, seed=1)
torch.cuda.synchronize()
print("ok", float(x.abs().max()))
