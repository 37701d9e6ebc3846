"""VAE decode time against the number of images per call (each count is one plan of that size).
usage: python scripts/decode_sweep.py [n ...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "diffusion-model_b200")]
import torch
from models.vae import VAE
from dmb200 import context_for
dev = torch.device("cuda", 0)
torch.manual_seed(0)
vae = VAE().eval().to(dev)
ctx = context_for(dev, os.environ.get("DMB200_PRECISION", "bf16")); ctx.ensure_vae(vae)
for n in [int(a) for a in sys.argv[1:]] or [2, 4, 8, 16, 32]:
    z = torch.randn((n, 4, 28, 28), device=dev)
    for _ in range(3): ctx.vae_decode(z, want_f32=False, want_u8=True)
    torch.cuda.synchronize()
    reps = max(4, 256 // n)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): ctx.vae_decode(z, want_f32=False, want_u8=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"n={n:3d}: {ms:.3f} ms per call, {ms / n * 32:.3f} ms per 32 images", flush=True)
