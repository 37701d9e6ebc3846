"""The oracle restatement (oracle/ref_torch.py) against vectors produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import ref_torch as R
from oracle import synth

TAPS = ["inc", "down1", "sa1", "down2", "sa2", "down3", "sa3", "bot1", "bot2", "bot3",
        "up1", "sa4", "up2", "sa5", "up3", "sa6"]


def _t(a):
    return torch.from_numpy(np.asarray(a))


def test_torch_version_matches_golden(golden):
    # the oracle is "reference source x this torch build" 
    assert str(golden["torch_version"]) == torch.__version__


def test_product_init_matches_reference_weights(golden, seeded_models):
    unet, vae = seeded_models
    sd, vsd = unet.state_dict(), vae.state_dict()
    assert list(sd) == [str(k) for k in golden["unet_keys"]]
    assert list(vsd) == [str(k) for k in golden["vae_keys"]]
    assert len(sd) == 191 and len(vsd) == 54
    np.testing.assert_allclose([float(v.double().sum()) for v in sd.values()], golden["unet_sums"], rtol=0, atol=0)
    np.testing.assert_allclose([float(v.double().sum()) for v in vsd.values()], golden["vae_sums"], rtol=0, atol=0)
    import hashlib
    h = hashlib.sha256()
    for v in sd.values():
        h.update(v.numpy().tobytes())
    assert h.hexdigest() == str(golden["unet_sha256"])


def test_unet_forward_and_taps(golden, seeded_models):
    sd = seeded_models[0].state_dict()
    taps = {}
    with torch.no_grad():
        eps, geom = R.unet_forward(sd, _t(golden["A_x"]), _t(golden["A_t"]), _t(golden["A_y"]),
                                   _t(golden["A_vals"]), _t(golden["A_mask"]), taps=taps)
        eps_nc, _ = R.unet_forward(sd, _t(golden["A_x"]), _t(golden["A_t"]), _t(golden["A_y"]))
    # explicit attention math vs torch's fused MHA kernel: same arithmetic, different summation order
    np.testing.assert_allclose(eps.numpy(), golden["A_eps"], rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(eps_nc.numpy(), golden["A_eps_nocond"], rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(geom.numpy(), golden["A_geom"], rtol=2e-4, atol=2e-5)
    for n in TAPS:
        assert tuple(taps[n].shape) == tuple(golden[f"A_tap_{n}_shape"])
        flat = taps[n].numpy().reshape(-1)
        scale = float(golden[f"A_tap_{n}_stat"][2])
        np.testing.assert_allclose(flat[golden[f"A_tap_{n}_idx"]], golden[f"A_tap_{n}_val"], rtol=0,
                                   atol=2e-5 * max(scale, 1.0), err_msg=n)


@pytest.mark.parametrize("tt", [1000, 500, 1])
def test_teacher_forced_step(golden, seeded_models, tt):
    sd = seeded_models[0].state_dict()
    _, alphas, alpha_bars = R.schedule(1000)
    t = torch.full((2,), tt, dtype=torch.long)
    with torch.no_grad():
        xo, _ = R.denoise_cond(sd, _t(golden["B_x"]), t, _t(golden["A_y"]), 3.0, 0, _t(golden["A_vals"]),
                               _t(golden["A_mask"]), _t(golden[f"B_noise_{tt}"]), alphas, alpha_bars)
    ref = golden[f"B_xout_{tt}"]
    np.testing.assert_allclose(xo.numpy(), ref, rtol=0, atol=5e-5 * np.abs(ref).max())


def test_short_trajectory_decode_and_metrics(golden, seeded_models):
    sd, vsd = seeded_models[0].state_dict(), seeded_models[1].state_dict()
    T = 20
    x_T, noise = synth.noise_bank(T, (2, 4, 28, 28), seed=1234)
    y = torch.tensor([1, 3])
    with torch.no_grad():
        lat = R.sample_latent(sd, x_T, y, _t(golden["A_vals"]), _t(golden["A_mask"]),
                              lambda i: noise[T - i], num_timesteps=T)
        img = R.vae_decode(vsd, lat)
    ref = golden["C_latent"]
    np.testing.assert_allclose(lat.numpy(), ref, rtol=0, atol=2e-4 * np.abs(ref).max())
    np.testing.assert_allclose(img.numpy().reshape(-1)[golden["C_img_idx"]], golden["C_img_val"], atol=2e-4)
    u8 = R.to_u8(img).numpy()
    assert (np.abs(u8.astype(int) - golden["C_img_u8"].astype(int)) > 1).sum() == 0
    assert (u8 != golden["C_img_u8"]).mean() < 1e-3
    masks = np.stack([R.binarise(golden["C_img_u8"][i]) for i in range(2)])
    assert (masks == golden["C_masks"]).all()  # Pillow's RGB->L restated exactly
    m = R.compute_metrics(golden["C_masks"][0], golden["C_masks"][1])
    for k, v in zip(golden["C_metric_keys"], golden["C_metric_vals"]):
        assert m[str(k)] == pytest.approx(float(v), rel=1e-12, abs=1e-12), k


def test_vae_decode(golden, seeded_models):
    vsd = seeded_models[1].state_dict()
    with torch.no_grad():
        im = R.vae_decode(vsd, _t(golden["D_z"]))
    np.testing.assert_allclose(im.numpy().reshape(-1)[golden["D_idx"]], golden["D_val"], atol=1e-5)
    assert (R.to_u8(im).numpy() != golden["D_u8"]).mean() < 1e-4


@pytest.mark.parametrize("cls", [1, 2, 3])
def test_condition_frontend(golden, cls):
    v, m = R.build_vals_mask(golden["E_rows"], cls)
    assert v.dtype == np.float32
    np.testing.assert_array_equal(v, golden[f"E_vals_{cls}"])
    np.testing.assert_array_equal(m, golden[f"E_mask_{cls}"])


def test_schedule(golden):
    betas, alphas, abars = R.schedule(1000)
    np.testing.assert_array_equal(betas.numpy(), golden["F_betas"])
    np.testing.assert_array_equal(abars.numpy(), golden["F_alpha_bars"])


def test_metrics_edge_cases():
    z = np.zeros((8, 8), bool)
    o = z.copy(); o[2, 2:5] = True
    assert R.compute_metrics(z, z)["iou"] == 1.0 and R.compute_metrics(z, z)["gauss_recall"] == 1.0
    m = R.compute_metrics(o, z)
    assert m["iou"] == 0.0 and m["far_noise_ratio"] == 0.0 and m["gauss_recall"] == 0.0
    with pytest.raises(ValueError):
        R.compute_metrics(z, np.zeros((4, 4), bool))


def test_scoring_oracle_matches_scipy():
    """Pins the two third-party behaviours the on-GPU scorer (csrc/score.cuh) restates: scipy's
    exact Euclidean distance transform -- including its answer for an input with no background
    pixel, sqrt((y+1)^2 + x^2) (a virtual background pixel at (-1, 0)), which far_noise_ratio reaches when the ground truth is
    empty (eval_iou_noise.py:162-171, 211-232) -- and Pillow's integer RGB -> L conversion
    (eval_iou_noise.py:87)."""
    from scipy.ndimage import distance_transform_edt
    from PIL import Image
    yy, xx = np.mgrid[0:9, 0:13]
    d = distance_transform_edt(np.ones((9, 13), bool))
    assert np.array_equal(d, np.sqrt(((yy + 1) ** 2 + xx ** 2).astype(np.float64)))
    rng = np.random.default_rng(3)
    gt = rng.random((17, 23)) < 0.05
    gy, gx = np.nonzero(gt)
    y17, x23 = np.mgrid[0:17, 0:23]
    brute = ((y17[..., None] - gy) ** 2 + (x23[..., None] - gx) ** 2).min(-1)   # exact squared distances
    assert np.array_equal(distance_transform_edt(~gt), np.sqrt(brute.astype(np.float64)))
    rgb = rng.integers(0, 256, (3, 31, 29), dtype=np.uint8)
    pil = np.array(Image.fromarray(np.moveaxis(rgb, 0, -1), mode="RGB").convert("L"))
    assert np.array_equal(R.u8_to_gray(rgb), pil)


# ---- VAE.encode and Diffuser.add_noise (tests/golden/make_golden_encode.py)
@pytest.fixture(scope="module")
def enc_golden():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_encode.npz"))


def test_vae_encode_oracle_matches_reference(enc_golden, seeded_models):
    _, vae = seeded_models
    vsd = vae.state_dict()
    np.testing.assert_allclose([float(v.double().sum()) for v in vsd.values()], enc_golden["vae_sums"], rtol=0, atol=0)
    with torch.no_grad():
        z, kl, mu, logvar = R.vae_encode(vsd, _t(enc_golden["G_x"]).float(), _t(enc_golden["G_eps"]))
    # same torch build, same expressions: the restatement reproduces the reference bit for bit
    assert torch.equal(z, _t(enc_golden["G_z"]))
    assert torch.equal(mu, _t(enc_golden["G_mu"])) and torch.equal(logvar, _t(enc_golden["G_logvar"]))
    assert float(kl) == float(enc_golden["G_kl"])


def test_add_noise_oracle_matches_reference(enc_golden):
    _, _, alpha_bars = R.schedule(1000)
    x_t = R.add_noise(_t(enc_golden["H_x0"]), _t(enc_golden["H_t"]), _t(enc_golden["H_noise"]), alpha_bars)
    assert torch.equal(x_t, _t(enc_golden["H_xt"]))


def test_t1000_trajectory_fixture_pins_the_oracle_step(golden, seeded_models):
    """tests/golden/golden_t1000.npz (the unmodified reference's T = 1000 run, B = 2): the oracle's
    step function reproduces the recorded x_999 from x_1000 and x_1 -> x_0 (t = 1: no noise, the
    clamped alpha_bar_prev), teacher-forced from the reference's own states; the scorer reproduces
    its metrics from its masks."""
    import os
    G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_t1000.npz"))
    assert str(G["torch_version"]) == torch.__version__
    sd = seeded_models[0].state_dict()
    T = 1000
    x_T, noise = synth.noise_bank(T, (2, 4, 28, 28), seed=int(G["seed"]))
    rec = {int(t): _t(x) for t, x in zip(G["record_t"], G["record_x"])}
    assert torch.equal(rec[1000], x_T)
    _, alphas, abars = R.schedule(T)
    vals, mask, y = _t(golden["A_vals"]), _t(golden["A_mask"]), torch.tensor([1, 3])
    with torch.no_grad():
        for t_from, ref_next in ((1000, rec[999]), (2, rec[1]), (1, _t(G["latent"]))):
            t = torch.full((2,), t_from, dtype=torch.long)
            nxt, _ = R.denoise_cond(sd, rec[t_from], t, y, 3.0, 0, vals, mask, noise[T - t_from], alphas, abars)
            scale = float(ref_next.abs().max())
            np.testing.assert_allclose(nxt.numpy(), ref_next.numpy(), rtol=0, atol=2e-5 * scale, err_msg=str(t_from))
    m = R.compute_metrics(G["masks"][0], G["masks"][1])
    np.testing.assert_allclose([m[k] for k in sorted(m)], G["metric_vals"], rtol=1e-12)
    assert [str(k) for k in G["metric_keys"]] == sorted(m)
