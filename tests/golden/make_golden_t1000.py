"""Generate tests/golden/golden_t1000.npz: ONE full-length trajectory (T = 1000, B = 2) of the
UNMODIFIED reference (`Diffuser.sample_latent_cond`, diff.py:174-369, CPU fp32, torch of this
image) with injected noise, its decode, the binarised masks and `compute_metrics`.  Build
container only (needs /root/reference); about two minutes on 8 host threads.

    python tests/golden/make_golden_t1000.py

x_T and the per-step noise are NOT stored (25 MB): the tests regenerate them from
`oracle.synth.noise_bank(1000, (2,4,28,28), seed=4321)`.  The reference's own
`torch.randn_like(x)` call (diff.py:158) receives x_t BEFORE the step, so the hook that injects
the noise also records that state at a few timesteps: those pin `x_dump[k]`, k > 0
(generate_steps.py:160-178 saves exactly that state).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import load_reference, synth  # noqa: E402

T = 1000
SEED = 4321
RECORD_T = (1000, 999, 900, 750, 500, 250, 100, 10, 2, 1)   # x_t before the step at these t


def main():
    torch.set_num_threads(8)
    g0 = np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))
    vals, mask = torch.from_numpy(g0["A_vals"]), torch.from_numpy(g0["A_mask"])
    out = {}
    with load_reference.reference_modules() as ref:
        torch.manual_seed(0)
        unet = ref.unet_cond_geom.UnetCondWithGeomHead().eval()
        vae = ref.vae.VAE().eval()
        d = ref.diff.Diffuser(num_timesteps=T, device="cpu")
        x_T, noise = synth.noise_bank(T, (2, 4, 28, 28), seed=SEED)

        recorded = {}
        real_randn, real_like = torch.randn, torch.randn_like
        calls = [0]

        def fake_randn(*a, **k):
            return x_T.clone()

        def fake_like(x, **k):
            i = calls[0]
            calls[0] += 1
            t = T - i
            if t in RECORD_T:
                recorded[t] = x.detach().clone()
            return noise[i].clone()

        torch.randn, torch.randn_like = fake_randn, fake_like
        try:
            lat = d.sample_latent_cond(unet, class_counts=[(1, 1), (3, 1)], vae=None, progress=False,
                                       guidance_scale=3.0, cond=vals, cond_mask=mask, z_shape=(4, 28, 28))
        finally:
            torch.randn, torch.randn_like = real_randn, real_like
        assert calls[0] == T
        with torch.no_grad():
            img = vae.decode(lat)
        u8 = (img * 255).clamp(0, 255).to(torch.uint8)
        out.update(torch_version=np.array(torch.__version__), seed=np.array(SEED), latent=lat.numpy(),
                   img_u8=u8.numpy(), record_t=np.array(sorted(recorded, reverse=True)),
                   record_x=np.stack([recorded[t].numpy() for t in sorted(recorded, reverse=True)]))
        masks, grey = [], []
        for i in range(2):
            pil = d.reverse_to_img(img[i])
            arr = np.array(pil.convert("L"), dtype=np.uint8)
            grey.append(arr)
            masks.append(arr < 128)
        out["masks"] = np.stack(masks)
        out["grey"] = np.stack(grey)
        # threshold-adjacent pixels: |L - 128| <= 2
        out["near_threshold"] = np.array([int((np.abs(g.astype(int) - 128) <= 2).sum()) for g in grey])
        m = ref.eval_iou_noise.compute_metrics(masks[0], masks[1])
        out["metric_keys"] = np.array(sorted(m))
        out["metric_vals"] = np.array([m[k] for k in sorted(m)])
    path = os.path.join(ROOT, "tests", "golden", "golden_t1000.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; |latent|max", float(lat.abs().max()),
          "near-threshold pixels", out["near_threshold"].tolist(), "mask areas", [int(m.sum()) for m in masks])


if __name__ == "__main__":
    main()
