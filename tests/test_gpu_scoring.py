"""On-GPU scoring (dmb_score_masks, eval_iou_noise.py mirror) against the CPU oracle
(oracle/ref_torch.py::compute_metrics = the reference's compute_metrics with the scipy backend)
and the reference-generated golden metrics.  Bar: every integer-derived figure (iou, gt_iou,
far_noise_ratio, areas, fp) bit-identical; gauss_recall within 1e-12 relative (order of a
float64 sum)."""
import numpy as np
import pytest
import torch

from gpu_util import report
from oracle import ref_torch as R

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
EXACT = ("iou", "gt_iou", "far_noise_ratio", "inter", "union", "gt_area", "pred_area", "fp")


def _same(got, ref):
    for k in EXACT:
        assert got[k] == ref[k], (k, got, ref)
    assert abs(got["gauss_recall"] - ref["gauss_recall"]) <= 1e-12 * max(1.0, abs(ref["gauss_recall"])), (got, ref)


def _line_mask(rng, H, W, n_lines, jitter=0):
    m = np.zeros((H, W), bool)
    for _ in range(n_lines):
        y0, y1 = rng.integers(0, H, 2)
        x0, x1 = rng.integers(0, W, 2)
        n = 2 * max(H, W)
        ys = np.clip(np.round(np.linspace(y0, y1, n)).astype(int) + rng.integers(-jitter, jitter + 1), 0, H - 1)
        xs = np.clip(np.round(np.linspace(x0, x1, n)).astype(int) + rng.integers(-jitter, jitter + 1), 0, W - 1)
        m[ys, xs] = True
    return m


@pytest.mark.parametrize("H,W", [(224, 224), (7, 5), (254, 254), (33, 61), (1, 1), (64, 3)])
@pytest.mark.parametrize("sigma", [2.0, 1.5, 3.7])
def test_masks_match_oracle(H, W, sigma):
    import eval_iou_noise as E
    rng = np.random.default_rng(H * 1000 + W)
    pairs = []
    for i in range(6):
        gt = _line_mask(rng, H, W, 1 + i % 3)
        pred = _line_mask(rng, H, W, 1 + i % 3, jitter=2) | (rng.random((H, W)) < 0.01 * i)
        pairs.append((gt, pred))
    z = np.zeros((H, W), bool)
    pairs += [(z, pairs[0][1]), (pairs[0][0], z), (z, z), (~z, ~z), (pairs[1][0], pairs[1][0])]
    gt_t = torch.from_numpy(np.stack([p[0] for p in pairs]).astype(np.uint8)).to(DEV)
    pr_t = torch.from_numpy(np.stack([p[1] for p in pairs]).astype(np.uint8)).to(DEV)
    got = E.compute_metrics_batch(gt_t, pr_t, sigma=sigma)
    for (gt, pred), g in zip(pairs, got):
        _same(g, R.compute_metrics(gt, pred, sigma=sigma))


def test_rgb_images_are_binarised_like_pil():
    """(B,3,H,W) u8 images in, as they leave the VAE decode: PIL's integer luma + threshold."""
    import eval_iou_noise as E
    rng = np.random.default_rng(5)
    B, H, W = 5, 224, 224
    base = rng.integers(0, 256, (B, 1, H, W), dtype=np.uint8)
    pred = np.clip(base.astype(int) + rng.integers(-40, 41, (B, 3, H, W)), 0, 255).astype(np.uint8)
    gt = np.clip(base.astype(int) + rng.integers(-40, 41, (B, 3, H, W)), 0, 255).astype(np.uint8)
    for thr, inv in ((128, True), (100, True), (128, False)):
        got = E.compute_metrics_batch(torch.from_numpy(gt).to(DEV), torch.from_numpy(pred).to(DEV), threshold=thr,
                                      invert=inv)
        for b in range(B):
            gm, pm = R.u8_to_gray(gt[b]), R.u8_to_gray(pred[b])
            gm, pm = (gm < thr, pm < thr) if inv else (gm >= thr, pm >= thr)
            _same(got[b], R.compute_metrics(gm, pm))
    # grey (B,1,H,W) prediction against a (B,H,W) mask
    grey = rng.integers(0, 256, (B, 1, H, W), dtype=np.uint8)
    mask = rng.random((B, H, W)) < 0.03
    got = E.compute_metrics_batch(torch.from_numpy(mask.astype(np.uint8)).to(DEV), torch.from_numpy(grey).to(DEV))
    for b in range(B):
        _same(got[b], R.compute_metrics(mask[b], grey[b, 0] < 128))


def test_golden_masks_from_the_reference(golden):
    """The masks and metrics the unmodified reference produced (tests/golden/make_golden.py)."""
    import eval_iou_noise as E
    m = E.compute_metrics(golden["C_masks"][0].astype(bool), golden["C_masks"][1].astype(bool))
    ref = R.compute_metrics(golden["C_masks"][0].astype(bool), golden["C_masks"][1].astype(bool))
    _same(m, ref)
    report("scoring_golden", {k: m[k] for k in ("iou", "gt_iou", "far_noise_ratio", "gauss_recall")})
    want = dict(zip([str(k) for k in golden["C_metric_keys"]], [float(v) for v in golden["C_metric_vals"]]))
    for k in EXACT:
        assert m[k] == want[k], (k, m, want)
    assert abs(m["gauss_recall"] - want["gauss_recall"]) <= 1e-12


def test_single_mask_functions_and_errors():
    import eval_iou_noise as E
    from dmb200.native import NativeError
    z = np.zeros((8, 8), bool)
    o = z.copy(); o[2, 2:5] = True
    assert E.compute_metrics(z, z)["iou"] == 1.0 and E.compute_metrics(z, z)["gauss_recall"] == 1.0
    m = E.compute_metrics(o, z)
    assert m["iou"] == 0.0 and m["far_noise_ratio"] == 0.0 and m["gauss_recall"] == 0.0
    assert E.gaussian_weighted_recall(z, o) == 1.0 and E.far_noise_ratio(o, z) == 0.0
    assert E.far_noise_ratio(z, o) == R.compute_metrics(z, o)["far_noise_ratio"]   # scipy's empty-input answer
    with pytest.raises(ValueError):
        E.compute_metrics(z, np.zeros((4, 4), bool))
    with pytest.raises(ValueError):
        E.compute_metrics(o, o, sigma=0.0)
    with pytest.raises(NativeError):
        E.compute_metrics(np.zeros((300, 8), bool), np.zeros((300, 8), bool))


def test_scores_the_decoded_batch_in_place(golden):
    """Config 4's tail: decode on the device, score on the device, nothing but counters crosses PCIe."""
    import eval_iou_noise as E
    from models.vae import VAE
    torch.manual_seed(0)
    vae = VAE().eval().to(DEV)
    g = torch.Generator().manual_seed(3)
    z = torch.randn((8, 4, 28, 28), generator=g).to(DEV)
    u8 = vae.decode_u8(z)
    gt = torch.roll(u8, shifts=1, dims=0)
    got = E.compute_metrics_batch(gt, u8)
    u8_h, gt_h = u8.cpu().numpy(), gt.cpu().numpy()
    for b in range(8):
        _same(got[b], R.compute_metrics(R.binarise(gt_h[b]), R.binarise(u8_h[b])))


def test_command_line_run_directory(tmp_path, capsys):
    """eval_iou_noise.py's command line (eval_iou_noise.py:303-482) on the GPU scorer: ground truth
    p00000.jpg ... against pic1.png ..., one pair missing, two image sizes.  Every row of
    metrics_detail.csv must equal the oracle's compute_metrics on the masks the reference's
    load_binary_mask gives for the same files; the summary row and the run directory's layout are
    the reference's."""
    import pandas as pd
    from PIL import Image
    import eval_iou_noise as E
    rng = np.random.default_rng(11)
    gt_dir, gen_dir, out = tmp_path / "gt", tmp_path / "gen", tmp_path / "out"
    gt_dir.mkdir(); gen_dir.mkdir()
    sizes = [(224, 224)] * 5 + [(64, 96)] * 2
    for i, (H, W) in enumerate(sizes):
        gt = _line_mask(rng, H, W, 1 + i % 3)
        pred = _line_mask(rng, H, W, 1 + i % 3, jitter=1) | (rng.random((H, W)) < 0.004 * i)
        grey = np.where(gt, 20, 235).astype(np.uint8)                                  # dark strokes on white
        Image.fromarray(grey, mode="L").save(gt_dir / f"p{i:05d}.jpg", quality=92)      # (JPEG: ringing near the strokes)
        if i != 3:                                                                     # p00003 has no partner
            rgb = np.stack([np.where(pred, 30, 250)] * 3, -1).astype(np.uint8)
            rgb[..., 1] = np.clip(rgb[..., 1].astype(int) + rng.integers(-20, 20, (H, W)), 0, 255)
            Image.fromarray(rgb, mode="RGB").save(gen_dir / f"pic{i + 1}.png")
    run_dir = E.main(["--gt_dir", str(gt_dir), "--gen_dir", str(gen_dir), "--out_dir", str(out), "--invert",
                      "--sigma", "2.5", "--save_diff"], device=DEV)
    assert run_dir.parent == out and run_dir.name.startswith("run_")
    detail = pd.read_csv(run_dir / "metrics_detail.csv", encoding="utf-8-sig", float_precision="round_trip")
    summary = pd.read_csv(run_dir / "metrics_summary.csv", encoding="utf-8-sig", float_precision="round_trip")
    assert list(detail.columns) == ["gt_index", "gt_file", "gen_file", "gt_bin", "gen_bin", "pair_bin", "diff_bin", "iou",
                                    "gt_iou", "far_noise_ratio", "gauss_recall", "inter", "union", "gt_area", "pred_area", "fp"]
    assert list(detail["gt_index"]) == [0, 1, 2, 4, 5, 6]
    for _, row in detail.iterrows():
        gt = E.load_binary_mask(gt_dir / row["gt_file"], threshold=128, invert=True)
        pred = E.load_binary_mask(gen_dir / row["gen_file"], threshold=128, invert=True)
        _same({k: float(row[k]) for k in EXACT + ("gauss_recall",)}, R.compute_metrics(gt, pred, sigma=2.5))
        assert np.array_equal(np.array(Image.open(run_dir / row["gt_bin"])) == 255, gt)
        assert np.array_equal(np.array(Image.open(run_dir / row["gen_bin"])) == 255, pred)
        assert (run_dir / row["pair_bin"]).name == f"pair_gt{int(row['gt_index']):05d}_vs_pic{int(row['gt_index']) + 1}.png"
        assert (run_dir / row["diff_bin"]).is_file()
    s = summary.iloc[0]
    assert int(s["n_pairs"]) == 6 and int(s["missing_pairs_skipped"]) == 1 and float(s["sigma"]) == 2.5 and bool(s["invert"])
    assert abs(float(s["iou_mean"]) - detail["iou"].mean()) < 1e-12
    assert abs(float(s["gauss_recall_p90"]) - np.percentile(detail["gauss_recall"], 90)) < 1e-12
    cfg = dict(line.split("=", 1) for line in (run_dir / "config.txt").read_text().splitlines())
    assert list(cfg) == ["gt_dir", "gen_dir", "threshold", "invert", "sigma", "distance_backend", "max_pairs",
                         "save_diff", "missing_pairs_skipped"] and cfg["missing_pairs_skipped"] == "1"
    assert "[OK] run_dir:" in capsys.readouterr().out
    report("eval_cli", {"pairs": 6, "iou_mean": float(s["iou_mean"])})
