"""Parity of the VAE encoder and the forward diffusion step against the CPU
oracle (oracle/ref_torch.py) and the reference-generated vectors in golden_encode.npz.  All
calls go through the drop-in Python modules, i.e. through the C ABI.

Stated tolerances (bf16 mode = bf16 activations/weights, fp32 accumulate):
  posterior mean mu            rel-L2 <= 3e-2
  clamped logvar               max-abs <= 3e-2
  latent z (same eps)          rel-L2 <= 3e-2
  kl.mean()                    relative <= 3e-2
  add_noise                    bit-identical (fp32, every operation rounded separately)
"""
import os

import numpy as np
import pytest
import torch

from gpu_util import max_abs, rel_l2, report
from oracle import ref_torch as R

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_encode.npz")


@pytest.fixture(scope="module")
def enc_golden():
    return np.load(GOLDEN)


@pytest.fixture(scope="module")
def vae_pair():
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    torch.manual_seed(0)
    UnetCondWithGeomHead()
    vae = VAE().eval()
    vsd = {k: v.clone() for k, v in vae.state_dict().items()}
    return vae.to(DEV), vsd


def _t(a):
    return torch.from_numpy(np.asarray(a))


def test_encode_golden(enc_golden, vae_pair):
    vae, _ = vae_pair
    g = enc_golden
    x = _t(g["G_x"]).float().to(DEV)
    eps = _t(g["G_eps"]).to(DEV)
    z, kl, mu, logvar = vae.encode_moments(x, eps)
    rows = {"mu": rel_l2(mu, _t(g["G_mu"])), "logvar_max_abs": max_abs(logvar, _t(g["G_logvar"])),
            "z": rel_l2(z, _t(g["G_z"])), "kl_rel": abs(float(kl.mean()) - float(g["G_kl"])) / float(g["G_kl"])}
    report("vae_encode_golden", rows)
    assert rows["mu"] <= 3e-2 and rows["z"] <= 3e-2, rows
    assert rows["logvar_max_abs"] <= 3e-2, rows
    assert rows["kl_rel"] <= 3e-2, rows
    z2, klm = vae.encode(x, eps)
    assert torch.equal(z2, z) and abs(float(klm) - float(kl.mean())) < 1e-7


def test_encode_layers_match_oracle(enc_golden, vae_pair):
    """Every convolution output and plain GroupNorm output of the encoder against the oracle."""
    from dmb200 import context_for
    vae, vsd = vae_pair
    x = _t(enc_golden["G_x"]).float()
    eps = _t(enc_golden["G_eps"])
    taps = {}
    with torch.no_grad():
        R.vae_encode(vsd, x, eps, taps=taps)
    vae.encode(x.to(DEV), eps.to(DEV))
    ctx = context_for(DEV)
    rows = {}
    for name in ("venc.0", "venc.3", "venc.5", "venc.6", "venc.9", "venc.11", "venc.12", "venc.15", "venc.17"):
        got = ctx.get_tap(name)
        assert tuple(got.shape) == tuple(taps[name].shape), name
        rows[name] = rel_l2(got, taps[name])
    report("vae_encode_layers", rows)
    assert max(rows.values()) <= 3e-2, rows


@pytest.mark.parametrize("batch", [1, 5, 33])
def test_encode_matches_oracle(vae_pair, batch):
    """Seeded images against the oracle, incl. a batch that spans two chunks of 32."""
    vae, vsd = vae_pair
    g = torch.Generator().manual_seed(100 + batch)
    x = torch.rand((batch, 3, 224, 224), generator=g)
    x[:, :, 40:44, 20:200] = 0.0
    eps = torch.randn((batch, 4, 28, 28), generator=g)
    nref = min(batch, 3)                                 # the oracle on the first and last samples
    pick = list(range(nref - 1)) + [batch - 1] if batch > 1 else [0]
    with torch.no_grad():
        z_ref, _, mu_ref, lv_ref = R.vae_encode(vsd, x[pick], eps[pick])
    z, kl, mu, logvar = vae.encode_moments(x.to(DEV), eps.to(DEV))
    assert tuple(z.shape) == (batch, 4, 28, 28) and tuple(kl.shape) == (batch,)
    rows = {"mu": rel_l2(mu[pick], mu_ref), "z": rel_l2(z[pick], z_ref),
            "logvar_max_abs": max_abs(logvar[pick], lv_ref)}
    report(f"vae_encode_B{batch}", rows)
    assert rows["mu"] <= 3e-2 and rows["z"] <= 3e-2 and rows["logvar_max_abs"] <= 3e-2, rows
    # the head's own arithmetic, checked exactly on the device's mu / logvar
    mu_c, lv_c, eps_c = mu.cpu(), logvar.cpu(), eps
    z_from = (mu_c + eps_c * torch.exp(0.5 * lv_c)) * 0.18215
    assert float((z.cpu() - z_from).abs().max()) <= 1e-5 * float(z_from.abs().max())
    kl_from = 0.5 * torch.sum(torch.exp(lv_c.double()) + mu_c.double() ** 2 - 1.0 - lv_c.double(), dim=(1, 2, 3)) / (224 * 224)
    assert float((kl.cpu().double() - kl_from).abs().max()) <= 1e-5 * float(kl_from.abs().max())


def test_encode_is_per_sample(vae_pair):
    """A sample's latent does not depend on its neighbours in the batch (chunking, pad rows)."""
    vae, _ = vae_pair
    g = torch.Generator().manual_seed(5)
    x = torch.rand((4, 3, 224, 224), generator=g).to(DEV)
    eps = torch.randn((4, 4, 28, 28), generator=g).to(DEV)
    z4, _, _, _ = vae.encode_moments(x, eps)
    z1, _, _, _ = vae.encode_moments(x[2:3].contiguous(), eps[2:3].contiguous())
    assert torch.equal(z4[2:3], z1)


def test_encode_decode_round_trip_shapes(vae_pair):
    """encode -> decode runs on the same context; random-init weights, so only shape / range."""
    vae, _ = vae_pair
    x = torch.rand((2, 3, 224, 224), generator=torch.Generator().manual_seed(1)).to(DEV)
    z, _ = vae.encode(x)
    img = vae.decode(z)
    assert tuple(img.shape) == (2, 3, 224, 224)
    assert float(img.min()) >= 0.0 and float(img.max()) <= 1.0 and bool(torch.isfinite(img).all())


def test_encode_rejects_bad_input(vae_pair):
    vae, _ = vae_pair
    with pytest.raises(ValueError):
        vae.encode(torch.zeros((1, 3, 64, 64), device=DEV))
    with pytest.raises(ValueError):
        vae.encode(torch.zeros((1, 3, 224, 224), device=DEV), eps=torch.zeros((2, 4, 28, 28), device=DEV))


def test_add_noise_golden_bit_exact(enc_golden):
    from diff import Diffuser
    g = enc_golden
    d = Diffuser(num_timesteps=1000, device="cpu")     # the golden run's schedule (CPU cumprod)
    x_t, noise = d.add_noise(_t(g["H_x0"]).to(DEV), _t(g["H_t"]).to(DEV), noise=_t(g["H_noise"]).to(DEV))
    assert torch.equal(x_t.cpu(), _t(g["H_xt"]))
    assert torch.equal(noise.cpu(), _t(g["H_noise"]))


@pytest.mark.parametrize("shape", [(1, 4, 28, 28), (7, 4, 28, 28), (3, 3, 224, 224), (64, 5)])
def test_add_noise_matches_torch_bitwise(shape):
    """Against the reference's own expression evaluated by torch on the same device, and the
    default path that draws the noise from torch's generator like the reference."""
    from diff import Diffuser
    d = Diffuser(num_timesteps=1000, device=DEV)
    g = torch.Generator().manual_seed(3)
    x0 = torch.randn(shape, generator=g).to(DEV)
    t = torch.randint(1, 1001, (shape[0],), generator=g).to(DEV)
    t[0] = 1000
    torch.manual_seed(77)
    x_t, noise = d.add_noise(x0, t)
    torch.manual_seed(77)
    noise_ref = torch.randn_like(x0)
    assert torch.equal(noise, noise_ref)
    ab = d.alpha_bars[t - 1].view(-1, *([1] * (len(shape) - 1)))
    want = torch.sqrt(ab) * x0 + torch.sqrt(1 - ab) * noise_ref
    assert torch.equal(x_t, want)


def test_add_noise_asserts_range():
    from diff import Diffuser
    d = Diffuser(num_timesteps=1000, device=DEV)
    x0 = torch.zeros((2, 4, 28, 28), device=DEV)
    with pytest.raises(AssertionError):
        d.add_noise(x0, torch.tensor([0, 5], device=DEV))
    with pytest.raises(AssertionError):
        d.add_noise(x0, torch.tensor([5, 1001], device=DEV))


def test_encode_tf32_mode(enc_golden, vae_pair):
    """The fp32-storage / TF32-operand mode on the reference-generated vectors.
    Tolerances: mu / z rel-L2 <= 3e-3, logvar max-abs <= 3e-3."""
    from dmb200 import context_for
    vae, _ = vae_pair
    g = enc_golden
    ctx = context_for(DEV, "tf32")
    ctx.ensure_vae_enc(vae)
    z, kl, mu, logvar = ctx.vae_encode(_t(g["G_x"]).float().to(DEV), _t(g["G_eps"]).to(DEV), want_moments=True)
    rows = {"mu": rel_l2(mu, _t(g["G_mu"])), "logvar_max_abs": max_abs(logvar, _t(g["G_logvar"])),
            "z": rel_l2(z, _t(g["G_z"])), "kl_rel": abs(float(kl.mean()) - float(g["G_kl"])) / float(g["G_kl"])}
    report("vae_encode_golden_tf32", rows)
    assert rows["mu"] <= 3e-3 and rows["z"] <= 3e-3 and rows["logvar_max_abs"] <= 3e-3 and rows["kl_rel"] <= 3e-3, rows


def test_input_conv_paths_agree(enc_golden, vae_pair):
    """enc.0 on the CUDA cores (DMB200_ENC_IN_MMA=0, exact fp32 FFMA2) and on mma.sync TF32 (the
    default): both against the oracle's enc.0 output, tolerance 3e-3 (bf16 rounding of the output)."""
    from dmb200.native import Context
    vae, vsd = vae_pair
    x = _t(enc_golden["G_x"]).float()
    eps = _t(enc_golden["G_eps"])
    taps = {}
    with torch.no_grad():
        R.vae_encode(vsd, x, eps, taps=taps)
    rows = {}
    for flag in ("0", "1"):
        old = os.environ.get("DMB200_ENC_IN_MMA")
        os.environ["DMB200_ENC_IN_MMA"] = flag
        try:
            ctx = Context(DEV, "bf16")
        finally:
            if old is None:
                os.environ.pop("DMB200_ENC_IN_MMA", None)
            else:
                os.environ["DMB200_ENC_IN_MMA"] = old
        try:
            ctx.ensure_vae_enc(vae)
            ctx.vae_encode(x.to(DEV), eps.to(DEV))
            rows[f"mma{flag}"] = rel_l2(ctx.get_tap("venc.0"), taps["venc.0"])
        finally:
            ctx.close()
    report("vae_encode_input_conv", rows)
    assert max(rows.values()) <= 3e-3, rows


def test_vae_forward_matches_oracle(vae_pair):
    """VAE.forward, forward only (vae.py:71-76): (x_recon, z, loss, {'recon_mse', 'kl'}) against the oracle's
    encode -> decode -> F.mse_loss with the same normal draw; and the vae_recon_check.py sequence
    (vae_recon_check.py:25-31: encode, decode, per-image mse / mae).  Stated tolerances (bf16 mode):
    reconstruction max-abs <= 1.2e-2 (two networks in a row), recon_mse and kl relative <= 3e-2."""
    import torch.nn.functional as F
    vae, vsd = vae_pair
    g = torch.Generator().manual_seed(9)
    B = 5
    x = torch.rand((B, 3, 224, 224), generator=g)
    x = (x > 0.93).float() * 0.9 + 0.05 * torch.rand((B, 3, 224, 224), generator=g)   # sparse bright pixels on a dark field
    eps = torch.randn((B, 4, 28, 28), generator=g)
    z_ref, kl_ref, _, _ = R.vae_encode(vsd, x, eps)
    rec_ref = R.vae_decode(vsd, z_ref)
    mse_ref = F.mse_loss(rec_ref, x, reduction="mean")
    x_recon, z, loss, parts = vae(x.to(DEV), eps.to(DEV))
    assert set(parts) == {"recon_mse", "kl"} and loss.dim() == 0 and not loss.requires_grad
    m = max_abs(x_recon.cpu(), rec_ref)
    r_mse = abs(float(parts["recon_mse"]) - float(mse_ref)) / float(mse_ref)
    r_kl = abs(float(parts["kl"]) - float(kl_ref)) / abs(float(kl_ref))
    report("vae_forward", {"recon_max_abs": m, "mse_rel": r_mse, "kl_rel": r_kl, "z_rel_l2": rel_l2(z.cpu(), z_ref)})
    assert m <= 1.2e-2 and r_mse <= 3e-2 and r_kl <= 3e-2 and rel_l2(z.cpu(), z_ref) <= 3e-2
    assert abs(float(loss) - (float(parts["recon_mse"]) + 1e-6 * float(parts["kl"]))) <= 1e-6 * abs(float(loss))
    # the reduction itself: exact against float64 torch on the tensors the device produced, and repeatable
    from dmb200 import context_for
    ctx = context_for(DEV)
    sq, ab = ctx.recon_error(x_recon, x.to(DEV))
    d = (x_recon - x.to(DEV)).double()      # the difference in fp32, as the reference forms it; the sums in float64
    assert torch.allclose(sq, (d * d).flatten(1).sum(1), rtol=1e-12, atol=0)
    assert torch.allclose(ab, d.abs().flatten(1).sum(1), rtol=1e-12, atol=0)
    sq2, ab2 = ctx.recon_error(x_recon, x.to(DEV))
    assert torch.equal(sq, sq2) and torch.equal(ab, ab2)
    odd_a, odd_b = torch.rand((3, 1001), device=DEV), torch.rand((3, 1001), device=DEV)     # ragged: scalar path
    sq3, _ = ctx.recon_error(odd_a, odd_b)
    assert torch.allclose(sq3, ((odd_a - odd_b).double() ** 2).sum(1), rtol=1e-12, atol=0)
    # vae_recon_check.py:25-31 -- per-image figures from the same calls the script makes
    z_enc, _ = vae.encode(x.to(DEV), eps.to(DEV))
    x_rec = vae.decode(z_enc)
    mse = torch.mean((x_rec - x.to(DEV)) ** 2, dim=(1, 2, 3)).cpu()
    mse_o = torch.mean((rec_ref - x) ** 2, dim=(1, 2, 3))
    assert torch.allclose(mse, mse_o, rtol=3e-2)


def test_vae_forward_refuses_a_training_step(vae_pair):
    vae, _ = vae_pair
    x = torch.rand((1, 3, 224, 224), device=DEV)
    vae.train()
    try:
        with pytest.raises(NotImplementedError):
            vae(x)
        with torch.no_grad():
            assert vae(x)[0].shape == (1, 3, 224, 224)
    finally:
        vae.eval()
