"""Per-kernel device-time breakdown of one eager CFG step and one VAE decode
(dmb_set_kernel_timing).  usage: python scripts/profile_step.py [B] [precision]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "diffusion-model_b200")]
import torch
from oracle import synth
from models.unet_cond_geom import UnetCondWithGeomHead
from models.vae import VAE
from diff import Diffuser
from dmb200 import context_for

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
dev = torch.device("cuda", 0)
torch.manual_seed(0)
unet = UnetCondWithGeomHead().eval().to(dev); vae = VAE().eval().to(dev)
ctx = context_for(dev, prec); ctx.ensure_unet(unet); ctx.ensure_vae(vae)
coef = Diffuser(device=dev).step_coefficients()
y, vals, mask, _ = synth.mixed_conditions(B, 0)
y, vals, mask = y.to(dev), vals.to(dev), mask.to(dev)
x = torch.randn((B, 4, 28, 28), device=dev)
t = torch.full((B,), 500, dtype=torch.long, device=dev)
for _ in range(3):
    ctx.denoise_cond_step(x, t, y, 0, vals, mask, 3.0, coef, seed=1)
torch.cuda.synchronize()
ctx.set_kernel_timing(True)
reps = 5
for _ in range(reps):
    ctx.denoise_cond_step(x, t, y, 0, vals, mask, 3.0, coef, seed=1)
rows = ctx.kernel_timing()
ctx.set_kernel_timing(False)
tot = sum(r["ms"] for r in rows) / reps
print(f"== CFG step B={B} {prec}: {tot:.3f} ms (sum of per-kernel event times)")
for r in sorted(rows, key=lambda r: -r["ms"]):
    ms = r["ms"] / reps
    per = r["ms"] / r["count"]
    tf = r["flops"] / r["count"] / (per * 1e-3) / 1e12 if r["flops"] else 0
    gb = r["bytes"] / r["count"] / (per * 1e-3) / 1e9 if r["bytes"] else 0
    print(f"{r['label']:34s} n/step={r['count']//reps:3d} {ms:8.3f} ms/step {100*ms/tot:5.1f}%  per-launch {per*1000:8.1f} us  {tf:7.1f} TF/s {gb:7.0f} GB/s")
out = {"B": B, "precision": prec, "step_ms": tot, "rows": rows, "reps": reps}
z = torch.randn((min(B, 32), 4, 28, 28), device=dev)
ctx.vae_decode(z, False, True); torch.cuda.synchronize()
ctx.set_kernel_timing(True)
ctx.vae_decode(z, False, True)
vr = ctx.kernel_timing(); ctx.set_kernel_timing(False)
vt = sum(r["ms"] for r in vr)
print(f"== VAE decode n={z.shape[0]}: {vt:.3f} ms")
for r in sorted(vr, key=lambda r: -r["ms"]):
    per = r["ms"] / r["count"]
    tf = r["flops"] / r["count"] / (per * 1e-3) / 1e12 if r["flops"] else 0
    gb = r["bytes"] / r["count"] / (per * 1e-3) / 1e9 if r["bytes"] else 0
    print(f"{r['label']:34s} n={r['count']:3d} {r['ms']:8.3f} ms {100*r['ms']/vt:5.1f}%  {tf:7.1f} TF/s {gb:7.0f} GB/s")
out["vae_rows"] = vr
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"kernel_timing_B{B}_{prec}.json"), "w"), indent=1)
