"""The call sequences of the reference's two sampling scripts, executed on the B200 with the
drop-in modules (J1, A12, A13).  The reference tree does not travel to the GPU
box, so the sequences are replayed statement by statement; tests/test_scripts_unchanged.py checks
(in the build container) that the unmodified scripts bind to these very classes and signatures.

  generate_cond.py:41-77   string device "cuda" -> Utils.loadModel(.pth) x2 -> Diffuser ->
                           EntityCsvSampler -> .sample(csv, count) x3 with set_class_id ->
                           Utils.saveImages -> pic{i}.png
  generate_steps.py:97-189 torch.device -> Diffuser -> EntityCsvSampler(..., device=) ->
                           load_cond(count=1, start=row) -> per step: vae.decode(x) (fp32),
                           .clamp, reverse_to_img, pixel/t{i}.png, latent/ch{c:02d}/t{i}.png,
                           diffuser.denoise_cond(model, x, t, y=..., ...) called from Python
"""
import os

import numpy as np
import pytest
import torch
from PIL import Image

from gpu_util import max_abs, rel_l2, report
from oracle import ref_torch as R
from oracle import synth

pytestmark = pytest.mark.gpu


def _checkpoints(root, seed):
    """Random-init weights saved the way the reference saves them (utils.py:63-65)."""
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    from utils import Utils
    torch.manual_seed(seed)
    unet, vae = UnetCondWithGeomHead(), VAE()
    udir = os.path.join(root, "result", f"2026_01_27_08_0{seed}", "model_para")
    vdir = os.path.join(root, "vae", "2025_09_30_19_34")
    os.makedirs(udir, exist_ok=True)
    os.makedirs(vdir, exist_ok=True)
    Utils.saveModelParameter(udir, unet)
    vpath = os.path.join(vdir, "vae_best.pth")
    torch.save(vae.state_dict(), vpath)
    sd = {k: v.clone() for k, v in unet.state_dict().items()}
    vsd = {k: v.clone() for k, v in vae.state_dict().items()}
    return os.path.join(udir, "trained_para.pth"), vpath, sd, vsd


def _csvs(root, n):
    rows = synth.synth_rows(n, seed=0)
    paths = {}
    for name in ("line", "circle", "arc"):
        d = os.path.join(root, "data", f"{name}_224x224_test")
        os.makedirs(d, exist_ok=True)
        paths[name] = os.path.join(d, f"{name}_224x224_test.csv")
        synth.write_csv(paths[name], rows)
    return paths, rows


def test_generate_cond_call_sequence(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    from utils import Utils
    from diff import Diffuser
    from entityCsvSampler import EntityCsvSampler

    _checkpoints(".", 6)
    csv, _ = _csvs(".", 8)
    out_dir = os.path.join("./generated_by_cond", "lambda_0")
    dirs = {}
    for name in ("line", "circle", "arc"):
        dirs[name] = os.path.join(out_dir, name)
        os.makedirs(dirs[name], exist_ok=True)

    unet_cond_ckpt = "./result/2026_01_27_08_06/model_para/trained_para.pth"
    vae_ckpt = "./vae/2025_09_30_19_34/vae_best.pth"
    device = "cuda" if torch.cuda.is_available() else "cpu"          # generate_cond.py:41
    unet = Utils.loadModel(unet_cond_ckpt, UnetCondWithGeomHead(), device=device)
    vae = Utils.loadModel(vae_ckpt, VAE(), device=device)
    num_timesteps = 1000
    diffuser = Diffuser(num_timesteps=num_timesteps, device=device)
    image_count = 4                                                   # the script: 500
    sampler = EntityCsvSampler(diffuser=diffuser, model=unet, vae=vae, class_id=1)

    images = {}
    torch.manual_seed(123)
    images["line"] = sampler.sample(csv["line"], count=image_count)
    sampler.set_class_id(2)
    images["circle"] = sampler.sample(csv["circle"], count=image_count)
    sampler.set_class_id(3)
    images["arc"] = sampler.sample(csv["arc"], count=image_count)
    for name in ("line", "circle", "arc"):
        Utils.saveImages(dirs[name], images[name])

    # what came out: lists of PIL RGB 224x224, files pic1..pic4.png that reopen to the same pixels
    for name in ("line", "circle", "arc"):
        assert isinstance(images[name], list) and len(images[name]) == image_count
        for i, im in enumerate(images[name]):
            assert isinstance(im, Image.Image) and im.mode == "RGB" and im.size == (224, 224)
            back = np.array(Image.open(os.path.join(dirs[name], f"pic{i + 1}.png")))
            assert (back == np.array(im)).all()
    # the same global-generator seed reproduces the run: the tensor path returns the same pixels
    torch.manual_seed(123)
    sampler.set_class_id(1)
    vals, mask = sampler.load_cond(csv["line"], count=image_count)
    u8 = diffuser.sample_latent_cond(model=unet, class_counts=(1, vals.shape[0]), vae=vae, guidance_scale=3.0,
                                     cond=vals, cond_mask=mask, return_u8=True, progress=False)
    got = np.stack([np.array(im).transpose(2, 0, 1) for im in images["line"]])
    assert (u8.cpu().numpy() == got).all()
    assert len({im.tobytes() for n in images for im in images[n]}) == 3 * image_count   # all distinct


def _save_latent_channels_by_dir(z, step, latent_root):
    """generate_steps.py:36-66, restated: per-channel min-max -> u8 (truncation) -> ch{c:02d}/t{step}.png."""
    z = z[0].detach().cpu()
    for c in range(z.shape[0]):
        ch_dir = os.path.join(latent_root, f"ch{c:02d}")
        os.makedirs(ch_dir, exist_ok=True)
        ch = z[c]
        vmin, vmax = ch.min(), ch.max()
        ch_norm = (ch - vmin) / (vmax - vmin) if vmax > vmin else torch.zeros_like(ch)
        Image.fromarray((ch_norm.numpy() * 255).astype(np.uint8), mode="L").save(os.path.join(ch_dir, f"t{step}.png"))


def test_generate_steps_call_sequence_and_dumps(tmp_path, monkeypatch):
    """save_reverse_steps_for_csv_row (generate_steps.py:73-191) for T = 8, save_every = 1, with
    the noise of every torch.randn_like call injected so that the oracle can follow: every x_t
    BEFORE its step (what the script dumps) and every decoded pixel image against the oracle."""
    monkeypatch.chdir(tmp_path)
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    from utils import Utils
    from diff import Diffuser
    from entityCsvSampler import EntityCsvSampler

    upath, vpath, sd, vsd = _checkpoints(".", 6)
    csv, rows = _csvs(".", 60)
    device = "cuda"
    unet = Utils.loadModel(upath, UnetCondWithGeomHead(), device=device)
    vae = Utils.loadModel(vpath, VAE(), device=device)

    csv_path, row_index, class_id = csv["line"], 52, 1
    num_timesteps, z_shape, guidance_scale, null_label, base_wh = 8, (1, 4, 28, 28), 3.0, 0, (400, 400)
    x_T, noise = synth.noise_bank(num_timesteps, z_shape, seed=31)
    calls = [0]

    def fake_randn(*a, **k):
        return x_T.clone().to(k.get("device", "cpu"))

    def fake_like(x, **k):
        calls[0] += 1
        return noise[calls[0] - 1].clone().to(x.device)

    monkeypatch.setattr(torch, "randn", fake_randn)
    monkeypatch.setattr(torch, "randn_like", fake_like)

    # ---- the function body, generate_steps.py:97-189
    device_t = torch.device(device)
    B = z_shape[0]
    out_dir = os.path.join("./step_images", f"class_line_row{int(row_index):05d}")
    pixel_dir, latent_dir = os.path.join(out_dir, "pixel"), os.path.join(out_dir, "latent")
    os.makedirs(pixel_dir)
    os.makedirs(latent_dir)
    diffuser = Diffuser(num_timesteps=num_timesteps, device=device_t)
    sampler = EntityCsvSampler(diffuser=diffuser, model=unet, vae=vae, class_id=class_id, base_wh=base_wh,
                               device=device_t)
    vals, mask = sampler.load_cond(csv_path, count=1, start=row_index)
    y = torch.tensor([int(class_id)], device=device_t, dtype=torch.long)
    model_noise = unet.to(device_t)
    model_noise.eval()
    vae.eval()
    x = torch.randn(z_shape, device=device_t)
    save_set = set(range(num_timesteps, 0, -1))
    save_set.add(1)
    states, pixels = {}, {}
    for i in range(num_timesteps, 0, -1):
        if i in save_set:
            img = vae.decode(x)
            img = img.clamp(0, 1)
            pil = diffuser.reverse_to_img(img[0])
            pil.save(os.path.join(pixel_dir, f"t{i}.png"))
            _save_latent_channels_by_dir(z=x, step=i, latent_root=latent_dir)
            states[i], pixels[i] = x.detach().cpu().clone(), np.array(pil)
        t = torch.full((B,), i, device=device_t, dtype=torch.long)
        x = diffuser.denoise_cond(model_noise, x, t, y=y, guidance_scale=guidance_scale, null_label=null_label,
                                  cond_vals=vals, cond_mask=mask)
    assert calls[0] == num_timesteps
    monkeypatch.undo()
    monkeypatch.chdir(tmp_path)

    # ---- the oracle on the same row, same noise
    v_np, m_np = R.build_vals_mask(rows[row_index:row_index + 1], class_id)
    np.testing.assert_array_equal(vals.cpu().numpy(), v_np)
    record = []
    with torch.no_grad():
        x0_ref = R.sample_latent(sd, x_T, torch.tensor([class_id]), torch.from_numpy(v_np), torch.from_numpy(m_np),
                                 lambda i: noise[num_timesteps - i], num_timesteps=num_timesteps, record=record)
        before = {num_timesteps: x_T}
        for (i, _eps, x_after) in record:
            before[i - 1] = x_after
        errs, pix = {}, {}
        for i in range(num_timesteps, 0, -1):
            errs[i] = rel_l2(states[i], before[i])
            ref_u8 = R.to_u8(R.vae_decode(vsd, before[i]).clamp(0, 1))[0].numpy().transpose(1, 2, 0)
            pix[i] = int(np.abs(pixels[i].astype(int) - ref_u8.astype(int)).max())
            back = np.array(Image.open(os.path.join(pixel_dir, f"t{i}.png")))
            assert (back == pixels[i]).all()
            for c in range(4):
                assert os.path.isfile(os.path.join(latent_dir, f"ch{c:02d}", f"t{i}.png"))
    final = rel_l2(x.cpu(), x0_ref)
    report("generate_steps_T8", {"x_t_rel_l2": {str(k): v for k, v in errs.items()}, "pixel_u8_max_diff": pix,
                                 "final_rel_l2": final})
    assert errs[num_timesteps] == 0.0
    assert max(errs.values()) < 8e-3 and final < 8e-3, (errs, final)
    assert max(pix.values()) <= 3, pix


def test_load_model_replaces_the_native_weight_copy(tmp_path, monkeypatch):
    """Utils.loadModel (utils.py:68-73) into a module that has already run: the repacked device
    copy must follow the new checkpoint (A13), for the U-Net and the VAE; writes through `.data`
    (which bump no version counter) are picked up after `refresh_native()`."""
    monkeypatch.chdir(tmp_path)
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    from utils import Utils
    upath_a, vpath_a, sd_a, vsd_a = _checkpoints("a", 6)
    upath_b, vpath_b, sd_b, vsd_b = _checkpoints("b", 7)
    dev = "cuda"
    y, vals, mask, _ = synth.mixed_conditions(3, seed=9)
    g = torch.Generator().manual_seed(3)
    x = torch.randn((3, 4, 28, 28), generator=g)
    t = torch.tensor([900, 300, 2])
    z = torch.randn((1, 4, 28, 28), generator=g)

    unet = Utils.loadModel(upath_a, UnetCondWithGeomHead(), device=dev)
    vae = Utils.loadModel(vpath_a, VAE(), device=dev)
    with torch.no_grad():
        ref = {"a": R.unet_forward(sd_a, x, t, y, vals, mask)[0], "b": R.unet_forward(sd_b, x, t, y, vals, mask)[0]}
        vref = {"a": R.vae_decode(vsd_a, z), "b": R.vae_decode(vsd_b, z)}

    def run():
        eps, _ = unet(x.to(dev), t.to(dev), y.to(dev), cond_vals=vals.to(dev), cond_mask=mask.to(dev))
        return eps.cpu(), vae.decode(z.to(dev)).cpu()

    eps_a, img_a = run()
    assert rel_l2(eps_a, ref["a"]) < 1.3e-2 and max_abs(img_a, vref["a"]) < 8e-3   # measured 7.2e-3 / 5.4e-3
    assert rel_l2(eps_a, ref["b"]) > 0.5                      # the two weight sets are unrelated
    unet2 = Utils.loadModel(upath_b, unet, device=dev)       # same module objects, new checkpoint
    vae2 = Utils.loadModel(vpath_b, vae, device=dev)
    assert unet2 is unet and vae2 is vae
    eps_b, img_b = run()
    rows = {"eps_a": rel_l2(eps_a, ref["a"]), "eps_b": rel_l2(eps_b, ref["b"]),
            "img_a": max_abs(img_a, vref["a"]), "img_b": max_abs(img_b, vref["b"])}
    report("load_model_reload", rows)
    assert rows["eps_b"] < 1.3e-2 and rows["img_b"] < 8e-3, rows   # measured 8.4e-3 / 5.1e-3
    # load_state_dict straight on the module (what loadModel does inside) back to A
    unet.load_state_dict({k: v.to(dev) for k, v in sd_a.items()})
    eps_a2, _ = run()
    assert torch.equal(eps_a2, eps_a)
    # a write through .data is invisible to version counters: explicit refresh
    for k, p in unet.state_dict(keep_vars=True).items():
        p.data.copy_(sd_b[k])
    unet.refresh_native()
    eps_b2, _ = run()
    assert torch.equal(eps_b2, eps_b)
