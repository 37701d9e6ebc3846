"""Host-side logic of the drop-in modules (no GPU): schedule table, condition building,
CSV front-end, error behaviour, and that nothing runs on the CPU."""
import os

import numpy as np
import pytest
import torch

from oracle import load_reference
from oracle import ref_torch as R
from oracle import synth


def test_schedule_is_bit_identical(golden):
    from diff import Diffuser
    d = Diffuser(num_timesteps=1000)
    np.testing.assert_array_equal(d.betas.numpy(), golden["F_betas"])
    np.testing.assert_array_equal(d.alpha_bars.numpy(), golden["F_alpha_bars"])
    coef = d.step_coefficients()
    assert coef.shape == (1000, 3) and coef.dtype == torch.float32
    # same expressions as diff.py:141-144, 160-161
    t = torch.tensor([1, 2, 500, 1000])
    a, ab = d.alphas[t - 1], d.alpha_bars[t - 1]
    abp = d.alpha_bars[torch.clamp(t - 2, min=0)]
    assert torch.equal(coef[t - 1, 0], (1 - a) / torch.sqrt(1 - ab))
    assert torch.equal(coef[t - 1, 1], torch.sqrt(a))
    assert torch.equal(coef[t - 1, 2], torch.sqrt((1 - a) * (1 - abp) / (1 - ab)))


def test_step_update_formula_matches_oracle():
    """(x - c1 eps)/sqrt(a) + sigma noise with the table == the reference expression."""
    from diff import Diffuser
    d = Diffuser(num_timesteps=1000)
    _, alphas, abars = R.schedule(1000)
    g = torch.Generator().manual_seed(0)
    x, eps, nz = (torch.randn((3, 4, 28, 28), generator=g) for _ in range(3))
    for tt in (1000, 17, 1):
        t = torch.full((3,), tt, dtype=torch.long)
        ref = R.ddpm_update(x, eps, t, nz, alphas, abars)
        c = d.step_coefficients()[tt - 1]
        got = (x - c[0] * eps) / c[1] + (0 if tt == 1 else 1) * nz * c[2]
        assert torch.allclose(got, ref, rtol=0, atol=1e-6 * float(ref.abs().max()))


@pytest.mark.parametrize("cls", [1, 2, 3])
def test_csv_sampler_condition(tmp_path, golden, cls):
    from diff import Diffuser
    from entityCsvSampler import EntityCsvSampler
    path = os.path.join(tmp_path, "rows.csv")
    np.savetxt(path, golden["E_rows"], delimiter=",", fmt="%.6f")
    s = EntityCsvSampler(Diffuser(device="cpu"), None, None, class_id=cls, device=torch.device("cpu"))
    vals, mask = s.load_cond(path)
    np.testing.assert_array_equal(vals.numpy(), golden[f"E_vals_{cls}"])
    np.testing.assert_array_equal(mask.numpy(), golden[f"E_mask_{cls}"])
    v2, m2 = s.load_cond(path, count=3, start=5)
    np.testing.assert_array_equal(v2.numpy(), golden[f"E_vals_{cls}"][5:8])
    with pytest.raises(ValueError):
        s.load_cond(path, count=1, start=99)  # empty selection
    s.set_class_id(7)
    with pytest.raises(ValueError):
        s.load_cond(path)


def test_infer_base_wh(tmp_path):
    import pandas as pd
    from diff import Diffuser
    from entityCsvSampler import EntityCsvSampler
    rows = synth.synth_rows(8, seed=1)
    rows[0, 1] = 399.2
    rows[1, 2] = 280.9
    rows[:, [2, 4]] = np.minimum(rows[:, [2, 4]], 280.9)
    s = EntityCsvSampler(Diffuser(), None, None, class_id=1, base_wh=None, device=torch.device("cpu"))
    assert s._infer_base_wh(pd.DataFrame(rows), 1) == (400.0, 280.0)


def _ref_condition(kwargs, class_counts):
    """What the reference's sample_latent_cond builds, captured at its first denoise call."""
    captured = {}

    class Stop(Exception):
        pass

    with load_reference.reference_modules() as ref:
        d = ref.diff.Diffuser(num_timesteps=2, device="cpu")

        def fake(**kw):
            captured.update(kw)
            raise Stop()

        d.denoise_cond = fake
        try:
            d.sample_latent_cond(None, class_counts, z_shape=(4, 28, 28), progress=False, **kwargs)
        except Stop:
            pass
    return captured["y"], captured["cond_vals"], captured["cond_mask"]


@pytest.mark.skipif(not load_reference.available(), reason="reference tree not present")
@pytest.mark.parametrize("case", range(6))
def test_condition_building_matches_reference(case):
    from diff import _build_condition, _normalise_counts, KEY_ORDER, CLASS_KEYS
    cc = [{1: 2, 3: 1}, (2, 3), [(1, 1), (2, 1), (3, 1)], {3: 2}, (1, 2), [(2, 2)]][case]
    t = torch.arange(36, dtype=torch.float32).reshape(3, 12) / 36
    t[0, 3] = 0.0
    kw = [
        dict(cond={1: {"x1": .1, "y2": .9, "bogus": 1}, 3: {"ar": .2}}, cond_mask={1: {"x1": 0.0, "x2": 1.0}}),
        dict(cond=t),
        dict(cond=t, cond_mask=(t > 0.5).float()),
        dict(cond=[{"ax": .5}, {"theta1": .25, "theta2": .75}], cond_mask=[{"ax": 0.0}, {"ay": 1.0}]),
        dict(),
        dict(cond_mask=torch.ones(2, 12)),
    ][case]
    y_ref, v_ref, m_ref = _ref_condition(kw, cc)
    items = _normalise_counts(cc)
    y_list = [c for c, n in items for _ in range(n)]
    v, m = _build_condition(kw.get("cond"), kw.get("cond_mask"), y_list, list(KEY_ORDER), CLASS_KEYS, "cpu")
    assert y_list == y_ref.tolist()
    assert torch.equal(v, v_ref) and torch.equal(m, m_ref)


def test_argument_errors_match_reference_types():
    from diff import Diffuser, _build_condition, _normalise_counts, KEY_ORDER, CLASS_KEYS
    with pytest.raises(ValueError):
        _normalise_counts("nope")
    with pytest.raises(ValueError):
        _normalise_counts({1: 0})
    with pytest.raises(ValueError):
        _build_condition(torch.zeros(2, 11), None, [1, 1], list(KEY_ORDER), CLASS_KEYS, "cpu")
    with pytest.raises(ValueError):
        _build_condition(torch.zeros(2, 12), torch.zeros(2, 3), [1, 1], list(KEY_ORDER), CLASS_KEYS, "cpu")
    with pytest.raises(ValueError):
        _build_condition([{}], None, [1, 1], list(KEY_ORDER), CLASS_KEYS, "cpu")
    d = Diffuser(num_timesteps=10)
    x = torch.zeros(1, 4, 28, 28)
    with pytest.raises(AssertionError):
        d.denoise_cond(None, x, torch.tensor([11]), y=torch.tensor([1]), guidance_scale=3.0)
    with pytest.raises(UnboundLocalError):
        d.denoise_cond(None, x, torch.tensor([3]), y=torch.tensor([1]), guidance_scale=0.0)
    with pytest.raises(ValueError):
        d.sample_latent_cond(None, (1, 1), z_shape=None, vae=None)


def test_no_cpu_fallback(seeded_models):
    """The product fails loudly instead of computing on the CPU."""
    from dmb200 import NativeError
    unet, vae = seeded_models
    x = torch.zeros(1, 4, 28, 28)
    with pytest.raises(NativeError):
        unet(x, torch.tensor([1]), torch.tensor([1]))
    with pytest.raises(NativeError):
        vae.decode(x)
    with pytest.raises(RuntimeError):
        unet.inc(x)  # parameter holders do not compute


def test_no_cpu_fallback_for_encode_and_add_noise(seeded_models):
    """The N3 extension follows the same rule; argument checks still come first, as in the reference."""
    from dmb200 import NativeError
    from diff import Diffuser
    _, vae = seeded_models
    with pytest.raises(NativeError):
        vae.encode(torch.zeros(1, 3, 224, 224))
    with pytest.raises(ValueError):
        vae.encode(torch.zeros(1, 3, 32, 32))
    d = Diffuser(num_timesteps=1000, device="cpu")
    x0 = torch.zeros(2, 4, 28, 28)
    with pytest.raises(NativeError):
        d.add_noise(x0, torch.tensor([1, 1000]))
    with pytest.raises(AssertionError):
        d.add_noise(x0, torch.tensor([0, 5]))


def test_product_does_not_import_oracle():
    import re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "diffusion-model_b200")
    for dp, _, fns in os.walk(root):
        for fn in fns:
            if fn.endswith(".py"):
                src = open(os.path.join(dp, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), fn


def test_reverse_to_img_truncates():
    from diff import Diffuser
    x = torch.tensor([0.0, 0.999 / 255, 1.0 / 255, 0.5, 1.0, 1.5, -0.2]).reshape(1, 1, 7).repeat(3, 1, 1)
    img = Diffuser().reverse_to_img(x)
    assert img.mode == "RGB" and img.size == (7, 1)
    assert list(np.array(img)[0, :, 0]) == [0, 0, 1, 127, 255, 255, 0]


def test_strided_coefficients_match_ddim_formula():
    """Extension N1: the (c0, c1, 0) rows reproduce the DDIM eta=0 update of the oracle."""
    import sys, os
    from oracle import ref_torch as R
    from diff import Diffuser
    d = Diffuser(num_timesteps=1000, device="cpu")
    coef, stride = d.strided_coefficients(20)
    assert stride == 50 and tuple(coef.shape) == (1000, 3)
    _, _, abars = R.schedule(1000)
    g = torch.Generator().manual_seed(3)
    x = torch.randn((2, 4, 28, 28), generator=g)
    eps = torch.randn((2, 4, 28, 28), generator=g)
    for t in (1000, 550, 50):
        c0, c1, c2 = (float(v) for v in coef[t - 1])
        assert c2 == 0.0
        got = (x - c0 * eps) / c1
        want = R.ddim_update(x, eps, t, t - stride, abars)
        assert float((got - want).abs().max()) < 2e-5 * float(want.abs().max())
    # rows that are never visited are harmless (divide by one)
    assert float(coef[998, 1]) == 1.0 and float(coef[998, 0]) == 0.0
    import pytest
    with pytest.raises(ValueError):
        d.strided_coefficients(7)


def test_png_batch_writer_round_trips_pixels(tmp_path):
    """Extension N4 (output path): the native parallel PNG writer behind Utils.saveImages(tensor).
    Every file must decode (PIL) to exactly the pixels that went in, under the reference's file
    names (utils.py:216-224), for RGB and grey, odd sizes and every compression level used."""
    import numpy as np
    import torch
    from PIL import Image
    from utils import Utils
    from dmb200 import write_png_batch
    rng = np.random.default_rng(0)
    imgs = torch.from_numpy(rng.integers(0, 256, (37, 3, 224, 224), dtype=np.uint8))
    imgs[0] = 255
    imgs[1] = 0
    Utils.saveImages(str(tmp_path), imgs)
    for i in range(imgs.shape[0]):
        im = Image.open(tmp_path / f"pic{i + 1}.png")
        assert im.mode == "RGB" and im.size == (224, 224)
        assert np.array_equal(np.moveaxis(np.array(im), -1, 0), imgs[i].numpy()), i
    # same pixels as the reference's own path (ToPILImage-equivalent -> image.save)
    ref_dir = tmp_path / "ref"
    ref_dir.mkdir()
    Utils.saveImages(str(ref_dir), [Image.fromarray(np.moveaxis(imgs[i].numpy(), 0, -1), mode="RGB") for i in range(3)])
    for i in range(3):
        assert np.array_equal(np.array(Image.open(ref_dir / f"pic{i + 1}.png")), np.array(Image.open(tmp_path / f"pic{i + 1}.png")))
    grey = torch.from_numpy(rng.integers(0, 256, (5, 28, 31), dtype=np.uint8))
    for level in (0, 1, 9):
        paths = [str(tmp_path / f"g{level}_{i}.png") for i in range(5)]
        write_png_batch(grey, paths, threads=2, compression_level=level)
        for i, p in enumerate(paths):
            im = Image.open(p)
            assert im.mode == "L" and np.array_equal(np.array(im), grey[i].numpy())
    Utils.saveStepImages(str(tmp_path), imgs[:2], [1000, 1])
    assert (tmp_path / "t1000.png").exists() and (tmp_path / "t1.png").exists()
    with pytest.raises(ValueError):
        write_png_batch(grey, ["only_one.png"])
    from dmb200 import NativeError
    with pytest.raises(NativeError):
        write_png_batch(grey[:1], [str(tmp_path / "no_such_dir" / "x.png")])


# ---- eval_iou_noise.py command line: host logic (pairing, summary row, visuals); the scoring itself is a GPU test ----

REF_SUMMARY_COLUMNS = ["n_pairs", "missing_pairs_skipped", "threshold", "invert", "sigma", "distance_backend",
                       "iou_mean", "iou_std", "gt_iou_mean", "gt_iou_std",
                       "far_noise_ratio_mean", "far_noise_ratio_std", "far_noise_ratio_median",
                       "far_noise_ratio_p90", "far_noise_ratio_p95", "gauss_recall_mean", "gauss_recall_std",
                       "gauss_recall_median", "gauss_recall_p90", "gauss_recall_p95", "gauss_overdraw_rate_gt1",
                       "run_dir"]   # eval_iou_noise.py:413-447, in order


def test_eval_cli_pairs_files_like_the_reference(tmp_path):
    """p00000.jpg <-> pic1.png, p00001.jpg <-> pic2.png ...; other names are ignored, case is not
    (eval_iou_noise.py:34-71, 346-361)."""
    import eval_iou_noise as E
    gt, gen = tmp_path / "gt", tmp_path / "gen"
    gt.mkdir(); gen.mkdir()
    for n in ["p00002.jpg", "p00000.jpg", "P00001.JPG", "p00007.jpg", "q00003.jpg", "p00004.png", "p5.jpg.bak"]:
        (gt / n).write_bytes(b"")
    for n in ["pic3.png", "pic1.png", "PIC2.PNG", "pic10.png", "pic_4.png", "pic5.jpg"]:
        (gen / n).write_bytes(b"")
    (gt / "p00009.jpg").mkdir()   # a directory with a matching name is not a file
    assert [i for i, _ in E.list_gt_files(gt)] == [0, 1, 2, 7]
    assert [i for i, _ in E.list_gen_files(gen)] == [1, 2, 3, 10]
    assert E._extract_gt_index("p00012.jpg") == 12 and E._extract_gt_index("p12.jpeg") is None
    assert E._extract_gen_index("pic7.png") == 7 and E._extract_gen_index("pic7.PNG") == 7
    pairs, missing = E.pair_files(gt, gen)
    assert [(i, a.name, b.name) for i, a, b in pairs] == [(0, "p00000.jpg", "pic1.png"), (1, "P00001.JPG", "PIC2.PNG"),
                                                           (2, "p00002.jpg", "pic3.png")]
    assert missing == 1   # p00007 has no pic8
    assert len(E.pair_files(gt, gen, max_pairs=2)[0]) == 2
    assert len(E.pair_files(gt, gen, max_pairs=0)[0]) == 3 and len(E.pair_files(gt, gen, max_pairs=-1)[0]) == 3
    if load_reference.available():
        with load_reference.reference_modules() as ref:
            ref_gt = [(i, p.name) for i, p in ref.eval_iou_noise.list_gt_files(gt)]
            ref_gen = [(i, p.name) for i, p in ref.eval_iou_noise.list_gen_files(gen)]
        assert ref_gt == [(i, p.name) for i, p in E.list_gt_files(gt)]
        assert ref_gen == [(i, p.name) for i, p in E.list_gen_files(gen)]


def test_eval_cli_summary_row_and_statistics(tmp_path):
    import eval_iou_noise as E
    rng = np.random.default_rng(3)
    detail = [{"iou": float(a), "gt_iou": float(b), "far_noise_ratio": float(c), "gauss_recall": float(d)}
              for a, b, c, d in rng.random((37, 4)) * [1, 1, 1, 1.3]]
    row = E.summarise(detail, missing=2, threshold=128, invert=True, sigma=2.0, run_dir=tmp_path)
    assert list(row) == REF_SUMMARY_COLUMNS
    gr = np.array([r["gauss_recall"] for r in detail])
    fn = np.array([r["far_noise_ratio"] for r in detail])
    assert row["n_pairs"] == 37 and row["missing_pairs_skipped"] == 2
    assert row["gauss_recall_mean"] == float(gr.mean()) and row["gauss_recall_std"] == float(gr.std(ddof=0))
    assert row["gauss_recall_p90"] == float(np.percentile(gr, 90)) and row["far_noise_ratio_median"] == float(np.percentile(fn, 50))
    assert row["gauss_overdraw_rate_gt1"] == float((gr > 1.0).mean())
    assert all(np.isnan(v) for v in E.mean_std(np.array([])))
    assert np.isnan(E.overdraw_rate(np.array([]))) and np.isnan(E.quantiles(np.array([]), [50])["p50"])
    if load_reference.available():
        with load_reference.reference_modules() as ref:
            r = ref.eval_iou_noise
            want = (r.mean_std(gr), r.quantiles(fn, [50, 90, 95]), r.overdraw_rate(gr, 1.0))
        assert want == (E.mean_std(gr), E.quantiles(fn, [50, 90, 95]), E.overdraw_rate(gr, 1.0))


def test_eval_cli_visuals(tmp_path):
    """binarised / side-by-side / difference images: same pixels as the reference writes (:97-156)."""
    from PIL import Image
    import eval_iou_noise as E
    rng = np.random.default_rng(4)
    gt, gen = rng.random((20, 30)) < 0.3, rng.random((20, 30)) < 0.3
    E.save_side_by_side(gt, gen, tmp_path / "pair.png")
    E.save_diff_visual(gt, gen, tmp_path / "diff.png")
    pair = np.array(Image.open(tmp_path / "pair.png"))
    assert pair.shape == (20, 60) and np.array_equal(pair[:, :30] == 255, gt) and np.array_equal(pair[:, 30:] == 255, gen)
    diff = np.array(Image.open(tmp_path / "diff.png"))
    assert (diff[gt & gen] == [0, 0, 0]).all() and (diff[gt & ~gen] == [0, 0, 255]).all()
    assert (diff[~gt & gen] == [255, 0, 0]).all() and (diff[~gt & ~gen] == [255, 255, 255]).all()
    small = rng.random((10, 15)) < 0.5
    E.save_side_by_side(gt, small, tmp_path / "pair2.png")           # other size: resized for display
    assert np.array(Image.open(tmp_path / "pair2.png")).shape == (20, 60)
    if load_reference.available():
        with load_reference.reference_modules() as ref:
            ref.eval_iou_noise.save_side_by_side(gt, gen, tmp_path / "rpair.png")
            ref.eval_iou_noise.save_diff_visual(gt, gen, tmp_path / "rdiff.png")
        assert np.array_equal(np.array(Image.open(tmp_path / "rpair.png")), pair)
        assert np.array_equal(np.array(Image.open(tmp_path / "rdiff.png")), diff)


def test_eval_cli_errors(tmp_path):
    import eval_iou_noise as E
    (tmp_path / "gt").mkdir()
    with pytest.raises(FileNotFoundError):
        E.main(["--gt_dir", str(tmp_path / "gt"), "--gen_dir", str(tmp_path / "nope"), "--out_dir", str(tmp_path / "o")])
    (tmp_path / "gen").mkdir()
    with pytest.raises(RuntimeError):   # no pairs (:363-369)
        E.main(["--gt_dir", str(tmp_path / "gt"), "--gen_dir", str(tmp_path / "gen"), "--out_dir", str(tmp_path / "o")])
