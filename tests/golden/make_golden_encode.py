"""Generate tests/golden/golden_encode.npz: VAE.encode and Diffuser.add_noise of the UNMODIFIED
reference (from /root/reference, CPU fp32, torch of this image) on seeded inputs (SURVEY.md 8f
N3).  The weights are the ones make_golden.py uses (torch.manual_seed(0), U-Net then VAE), so
tests/conftest.py's seeded models reproduce them.  Both reference functions draw their noise
inside; the draw is reproduced here by seeding torch's global generator the same way, and it
is stored so that the oracle and the CUDA path get it as an input.

    python tests/golden/make_golden_encode.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import load_reference  # noqa: E402


def main():
    torch.set_num_threads(8)
    out = {}
    with load_reference.reference_modules() as ref:
        torch.manual_seed(0)
        ref.unet_cond_geom.UnetCondWithGeomHead()          # same construction order as make_golden.py
        vae = ref.vae.VAE().eval()
        out["vae_sums"] = np.array([float(v.double().sum()) for v in vae.state_dict().values()])

        # ---- G: encode of two synthetic images (line-drawing-like: white page, dark strokes)
        g = torch.Generator().manual_seed(21)
        x = torch.ones((2, 3, 224, 224))
        for b in range(2):
            for _ in range(6):
                y0, x0 = (int(v) for v in torch.randint(8, 200, (2,), generator=g))
                ln = int(torch.randint(10, 80, (1,), generator=g))
                if torch.rand((), generator=g) < 0.5:
                    x[b, :, y0:y0 + 2, x0:min(223, x0 + ln)] = 0.0
                else:
                    x[b, :, y0:min(223, y0 + ln), x0:x0 + 2] = 0.0
        x = (x + 0.02 * torch.randn(x.shape, generator=g)).clamp(0, 1)
        x = x.half().float()          # stored as float16: round first so that the fixture IS the input
        with torch.no_grad():
            torch.manual_seed(1234)
            z, kl = vae.encode(x)
            torch.manual_seed(1234)
            eps = torch.randn((2, 4, 28, 28))      # the draw randn_like(std) made above
            h = vae.enc(x)
            mu = vae.to_mu(h)
            logvar = vae.to_logvar(h).clamp(-30.0, 20.0)
            z2 = (mu + eps * torch.exp(0.5 * logvar)) * vae.scale_factor
        assert torch.equal(z, z2), "the stored eps is not the draw encode() made"
        out.update(G_x=x.numpy().astype(np.float16), G_eps=eps.numpy(), G_z=z.numpy(), G_kl=np.array(float(kl)),
                   G_mu=mu.numpy(), G_logvar=logvar.numpy())

        # ---- H: add_noise at four timesteps incl. both ends
        d = ref.diff.Diffuser(num_timesteps=1000, device="cpu")
        g = torch.Generator().manual_seed(22)
        x0 = torch.randn((4, 4, 28, 28), generator=g)
        t = torch.tensor([1, 37, 500, 1000])
        torch.manual_seed(99)
        x_t, noise = d.add_noise(x0, t)
        out.update(H_x0=x0.numpy(), H_t=t.numpy(), H_noise=noise.numpy(), H_xt=x_t.numpy())

    path = os.path.join(ROOT, "tests", "golden", "golden_encode.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
