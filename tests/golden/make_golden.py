"""Generate tests/golden/golden.npz by running the UNMODIFIED reference (from /root/reference,
CPU fp32, torch of this image) on seeded inputs.  Runs in the build container only; the
fixture it writes is what pins the oracle restatement (oracle/ref_torch.py) on machines where
the reference is absent.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import load_reference, synth  # noqa: E402

TAPS = ["inc", "down1", "sa1", "down2", "sa2", "down3", "sa3", "bot1", "bot2", "bot3",
        "up1", "sa4", "up2", "sa5", "up3", "sa6"]


def sample_idx(n, k=64, seed=7):
    return np.random.default_rng(seed).choice(n, size=min(k, n), replace=False)


def main():
    torch.set_num_threads(8)
    out = {}
    with load_reference.reference_modules() as ref:
        torch.manual_seed(0)
        unet = ref.unet_cond_geom.UnetCondWithGeomHead().eval()
        vae = ref.vae.VAE().eval()
        sd, vsd = unet.state_dict(), vae.state_dict()
        out["torch_version"] = np.array(torch.__version__)
        out["unet_keys"] = np.array(list(sd))
        out["vae_keys"] = np.array(list(vsd))
        out["unet_sums"] = np.array([float(v.double().sum()) for v in sd.values()])
        out["vae_sums"] = np.array([float(v.double().sum()) for v in vsd.values()])
        h = hashlib.sha256()
        for v in sd.values():
            h.update(v.numpy().tobytes())
        out["unet_sha256"] = np.array(h.hexdigest())

        # ---- A: one forward, B=2, mixed classes, two different timesteps
        g = torch.Generator().manual_seed(11)
        x = torch.randn((2, 4, 28, 28), generator=g) * 3.0
        t = torch.tensor([1000, 37])
        rows = synth.synth_rows(4, seed=3)
        samp = ref.entityCsvSampler.EntityCsvSampler.__new__(ref.entityCsvSampler.EntityCsvSampler)
        import pandas as pd
        df = pd.DataFrame(rows)
        v1, m1 = samp._build_vals_mask_for(df, 1, (400, 400))
        v3, m3 = samp._build_vals_mask_for(df, 3, (400, 400))
        vals = torch.from_numpy(np.stack([v1[0], v3[1]]))
        mask = torch.from_numpy(np.stack([m1[0], m3[1]]))
        y = torch.tensor([1, 3])
        taps = {}
        hooks = [getattr(unet, n).register_forward_hook(lambda m, i, o, n=n: taps.__setitem__(n, o.detach()))
                 for n in TAPS]
        with torch.no_grad():
            eps, geom = unet(x, t, y, cond_vals=vals, cond_mask=mask)
            eps_nocond, _ = unet(x, t, y)
        for hk in hooks:
            hk.remove()
        out.update(A_x=x.numpy(), A_t=t.numpy(), A_y=y.numpy(), A_vals=vals.numpy(), A_mask=mask.numpy(),
                   A_eps=eps.numpy(), A_geom=geom.numpy(), A_eps_nocond=eps_nocond.numpy())
        # taps come from the last hooked call (the no-cond one); redo the cond call for taps
        taps.clear()
        hooks = [getattr(unet, n).register_forward_hook(lambda m, i, o, n=n: taps.__setitem__(n, o.detach()))
                 for n in TAPS]
        with torch.no_grad():
            unet(x, t, y, cond_vals=vals, cond_mask=mask)
        for hk in hooks:
            hk.remove()
        for n in TAPS:
            flat = taps[n].numpy().reshape(-1)
            idx = sample_idx(flat.size)
            out[f"A_tap_{n}_shape"] = np.array(taps[n].shape)
            out[f"A_tap_{n}_idx"] = idx
            out[f"A_tap_{n}_val"] = flat[idx]
            out[f"A_tap_{n}_stat"] = np.array([flat.mean(), flat.std(), np.abs(flat).max()])

        # ---- B: teacher-forced single steps at T=1000 (denoise_cond with injected noise)
        diffuser = ref.diff.Diffuser(num_timesteps=1000, device="cpu")
        g = torch.Generator().manual_seed(12)
        xs = torch.randn((2, 4, 28, 28), generator=g) * 2.0
        for tt in (1000, 500, 1):
            nz = torch.randn((2, 4, 28, 28), generator=g)
            tv = torch.full((2,), tt, dtype=torch.long)
            with load_reference.injected_noise(xs, [nz]):
                xo = diffuser.denoise_cond(unet, xs, tv, y=y, guidance_scale=3.0, null_label=0,
                                           cond_vals=vals, cond_mask=mask)
            out[f"B_noise_{tt}"] = nz.numpy()
            out[f"B_xout_{tt}"] = xo.numpy()
        out["B_x"] = xs.numpy()

        # ---- C: full sample_latent_cond + decode on a short schedule (T=20), B=2
        T = 20
        d20 = ref.diff.Diffuser(num_timesteps=T, device="cpu")
        x_T, noise = synth.noise_bank(T, (2, 4, 28, 28), seed=1234)
        with load_reference.injected_noise(x_T, list(noise)):
            lat = d20.sample_latent_cond(unet, class_counts=[(1, 1), (3, 1)], vae=None, progress=False,
                                         guidance_scale=3.0, cond=vals, cond_mask=mask, z_shape=(4, 28, 28))
        with torch.no_grad():
            img = vae.decode(lat)
        u8 = (img * 255).clamp(0, 255).to(torch.uint8)
        # x_T / noise are regenerated from synth.noise_bank(20, (2,4,28,28), seed=1234) by the tests
        out.update(C_latent=lat.numpy(), C_img_u8=u8.numpy())
        flat = img.numpy().reshape(-1)
        idx = sample_idx(flat.size, 4096)
        out.update(C_img_idx=idx, C_img_val=flat[idx])
        # mask + metrics through the reference scorer
        from PIL import Image
        masks = []
        for i in range(2):
            pil = d20.reverse_to_img(img[i])
            arr = np.array(pil.convert("L"), dtype=np.uint8)
            masks.append(arr < 128)
        out["C_masks"] = np.stack(masks)
        m = ref.eval_iou_noise.compute_metrics(masks[0], masks[1])
        out["C_metric_keys"] = np.array(sorted(m))
        out["C_metric_vals"] = np.array([m[k] for k in sorted(m)])

        # ---- D: VAE decode of a plain latent
        g = torch.Generator().manual_seed(13)
        z = torch.randn((1, 4, 28, 28), generator=g)
        with torch.no_grad():
            im = vae.decode(z)
        flat = im.numpy().reshape(-1)
        idx = sample_idx(flat.size, 4096, seed=8)
        out.update(D_z=z.numpy(), D_idx=idx, D_val=flat[idx],
                   D_u8=(im * 255).clamp(0, 255).to(torch.uint8).numpy())

        # ---- E: condition front-end on 16 synthetic rows, all three classes
        rows16 = synth.synth_rows(16, seed=0)
        rows16[3, 11] = 0.25   # angle already normalised -> passes through
        rows16[4, 12] = 725.0  # wraps mod 360
        df = pd.DataFrame(rows16)
        out["E_rows"] = rows16
        for cls in (1, 2, 3):
            v, mk = samp._build_vals_mask_for(df, cls, (400, 400))
            out[f"E_vals_{cls}"] = v
            out[f"E_mask_{cls}"] = mk

        # ---- F: schedule
        out["F_betas"] = diffuser.betas.numpy()
        out["F_alpha_bars"] = diffuser.alpha_bars.numpy()

    path = os.path.join(ROOT, "tests", "golden", "golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
