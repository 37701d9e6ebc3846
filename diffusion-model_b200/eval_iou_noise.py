"""Drop-in for the reference's eval_iou_noise.py -- its scoring functions computed on the B200
 and its command line (file pairing, run directory, CSV summary) around them.
Same names, arguments, return keys and exceptions as the reference:

  load_binary_mask(image_path, threshold=128, invert=True)     eval_iou_noise.py:77-94
  gaussian_weighted_recall(gt, pred, sigma=2.0)                eval_iou_noise.py:185-208
  far_noise_ratio(gt, pred, sigma=2.0)                         eval_iou_noise.py:211-232
  compute_metrics(gt, pred, sigma=2.0) -> dict                 eval_iou_noise.py:239-272

plus compute_metrics_batch(), which scores a whole batch of decoded u8 images that is already in
HBM (what the 4096-sample config needs: the reference's per-image scipy distance transform on
the host is the wall-clock tail there).  The device returns exact integer counters and one
float64 sum per image (dmb_score_masks); the ratios are formed here in float64 exactly as the
reference forms them, so iou / gt_iou / far_noise_ratio are bit-identical and gauss_recall
differs only by the order of a float64 sum.  There is no CPU implementation in this module.

  python eval_iou_noise.py --gt_dir G --gen_dir D --out_dir O [--threshold 128] [--invert]
                           [--sigma 2.0] [--max_pairs -1] [--save_diff]           eval_iou_noise.py:303-482
writes O/run_<timestamp>/{binarized/{gt,gen,pair}, diff, metrics_detail.csv, metrics_summary.csv,
config.txt} with the reference's file names and columns; all pairs are scored in one launch."""
from __future__ import annotations

from pathlib import Path
from typing import Dict, List

import numpy as np
import torch

from dmb200 import context_for


def load_binary_mask(image_path: Path, threshold: int = 128, invert: bool = True) -> np.ndarray:
    """File decoding stays on the host (PIL), like the reference; eval_iou_noise.py:77-94."""
    from PIL import Image
    img = Image.open(image_path).convert("L")
    arr = np.array(img, dtype=np.uint8)
    mask = arr < threshold if invert else arr >= threshold
    return mask.astype(bool)


def _device(device=None) -> torch.device:
    return torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())


def _metrics_from_counters(c, hit: float) -> Dict[str, float]:
    inter, union, gt_area, pred_area, fp, far = (int(v) for v in c)
    return {
        "iou": float(inter / union) if union > 0 else 1.0,                    # eval_iou_noise.py:250
        "gt_iou": float(inter / gt_area) if gt_area > 0 else 1.0,              # :253
        "far_noise_ratio": float(far / pred_area) if pred_area > 0 else 0.0,   # :221-232
        "gauss_recall": float(hit / gt_area) if gt_area > 0 else 1.0,          # :193-208
        "inter": float(inter), "union": float(union), "gt_area": float(gt_area),
        "pred_area": float(pred_area), "fp": float(fp),
    }


def compute_metrics_batch(gt: torch.Tensor, pred: torch.Tensor, sigma: float = 2.0, threshold: int = 128,
                          invert: bool = True) -> List[Dict[str, float]]:
    """gt / pred: uint8 CUDA tensors, (B,3,H,W) RGB or (B,1,H,W) grey images (binarised like
    load_binary_mask) or (B,H,W) masks.  One dict per image, keys as compute_metrics."""
    if sigma <= 0:
        raise ValueError("sigma must be > 0")
    ctx = context_for(pred.device)
    counts, hit = ctx.score_masks(pred, gt, threshold=threshold, invert=invert, sigma=sigma)
    counts, hit = counts.cpu().tolist(), hit.cpu().tolist()
    return [_metrics_from_counters(c, h) for c, h in zip(counts, hit)]


def _as_device_mask(m: np.ndarray, dev: torch.device) -> torch.Tensor:
    m = np.asarray(m)
    if m.ndim != 2:
        raise ValueError(f"masks must be 2-D, got shape {m.shape}")
    return torch.from_numpy(np.ascontiguousarray(m.astype(np.uint8))).to(dev)[None]


def compute_metrics(gt: np.ndarray, pred: np.ndarray, sigma: float = 2.0, device=None) -> Dict[str, float]:
    if gt.shape != pred.shape:
        raise ValueError(f"Shape mismatch: gt{gt.shape} vs pred{pred.shape}")
    dev = _device(device)
    return compute_metrics_batch(_as_device_mask(gt, dev), _as_device_mask(pred, dev), sigma=sigma)[0]


def gaussian_weighted_recall(gt: np.ndarray, pred: np.ndarray, sigma: float = 2.0, device=None) -> float:
    if int(np.asarray(gt).sum()) == 0:   # the reference returns before it looks at sigma (:193-195)
        return 1.0
    return compute_metrics(gt, pred, sigma=sigma, device=device)["gauss_recall"]


def far_noise_ratio(gt: np.ndarray, pred: np.ndarray, sigma: float = 2.0, device=None) -> float:
    if int(np.asarray(pred).sum()) == 0:   # (:221-223)
        return 0.0
    return compute_metrics(gt, pred, sigma=sigma, device=device)["far_noise_ratio"]


# ---------------------------------------------------------------------------
# Command line (eval_iou_noise.py:303-482): pair p00000.jpg <-> pic1.png, score, write the run
# directory.  The reference scores pair by pair on the host; here the files are decoded on host
# threads, the masks of ALL pairs of one image size go to the GPU as one batch, and
# dmb_score_masks scores them in one launch.  File names, CSV columns and config keys are the
# reference's, so whatever reads its run directories reads these.
# ---------------------------------------------------------------------------
import re
from concurrent.futures import ThreadPoolExecutor
from datetime import datetime
from typing import Optional, Tuple

_NAME_RULES = {"gt": re.compile(r"^p(\d+)\.jpg$", re.IGNORECASE),      # p00000.jpg -> 0   (:34, 38-42)
               "gen": re.compile(r"^pic(\d+)\.png$", re.IGNORECASE)}   # pic1.png   -> 1   (:35, 45-49)
DISTANCE_BACKEND = "dmb200-cuda-edt"   # what the summary's distance_backend column says (the reference: scipy / opencv)


def _index_of(kind: str, name: str) -> Optional[int]:
    m = _NAME_RULES[kind].match(name)
    return int(m.group(1)) if m else None


def _extract_gt_index(name: str) -> Optional[int]:
    return _index_of("gt", name)


def _extract_gen_index(name: str) -> Optional[int]:
    return _index_of("gen", name)


def _list_indexed(kind: str, folder: Path) -> List[Tuple[int, Path]]:
    found = [(_index_of(kind, p.name), p) for p in Path(folder).iterdir() if p.is_file()]
    return sorted(((i, p) for i, p in found if i is not None), key=lambda ip: ip[0])


def list_gt_files(gt_dir: Path) -> List[Tuple[int, Path]]:
    return _list_indexed("gt", gt_dir)       # eval_iou_noise.py:52-60


def list_gen_files(gen_dir: Path) -> List[Tuple[int, Path]]:
    return _list_indexed("gen", gen_dir)     # eval_iou_noise.py:63-71


def pair_files(gt_dir: Path, gen_dir: Path, max_pairs: int = -1) -> Tuple[List[Tuple[int, Path, Path]], int]:
    """Pairs (gt_index, gt_path, gen_path) with gen index = gt index + 1, and how many ground-truth files had
    no partner (eval_iou_noise.py:346-361)."""
    gen_by_index = dict(list_gen_files(gen_dir))
    pairs, missing = [], 0
    for gt_idx, gt_path in list_gt_files(gt_dir):
        partner = gen_by_index.get(gt_idx + 1)
        if partner is None:
            missing += 1
        else:
            pairs.append((gt_idx, gt_path, partner))
    if max_pairs is not None and max_pairs > 0:
        pairs = pairs[:max_pairs]
    return pairs, missing


def mask_to_pil(mask: np.ndarray):
    from PIL import Image
    return Image.fromarray(np.asarray(mask, dtype=np.uint8) * 255, mode="L")     # :97-100


def save_side_by_side(gt_mask: np.ndarray, gen_mask: np.ndarray, out_path: Path) -> None:
    """GT | GEN as one grey image (:103-119); a generated mask of another size is resized (nearest) for display."""
    from PIL import Image
    left, right = mask_to_pil(gt_mask), mask_to_pil(gen_mask)
    if right.size != left.size:
        right = right.resize(left.size, resample=Image.NEAREST)
    canvas = Image.new("L", (2 * left.width, left.height), color=0)
    for k, im in enumerate((left, right)):
        canvas.paste(im, (k * left.width, 0))
    canvas.save(out_path)


def save_diff_visual(gt_mask: np.ndarray, gen_mask: np.ndarray, out_path: Path) -> None:
    """White background, agreement black, missing strokes blue, extra strokes red (:122-156)."""
    from PIL import Image
    gt_mask, gen_mask = np.asarray(gt_mask, bool), np.asarray(gen_mask, bool)
    colour = np.array([[255, 255, 255],    # neither
                       [255, 0, 0],        # generated only
                       [0, 0, 255],        # ground truth only
                       [0, 0, 0]], np.uint8)   # both
    Image.fromarray(colour[2 * gt_mask.astype(np.uint8) + gen_mask.astype(np.uint8)], mode="RGB").save(out_path)


def mean_std(x: np.ndarray) -> Tuple[float, float]:
    x = np.asarray(x, np.float64)
    return (float(x.mean()), float(x.std(ddof=0))) if x.size else (float("nan"), float("nan"))     # :275-278


def quantiles(x: np.ndarray, ps: List[float]) -> Dict[str, float]:
    x = np.asarray(x, np.float64)
    vals = np.percentile(x, ps) if x.size else [float("nan")] * len(ps)                              # :281-288
    return {f"p{int(p)}": float(v) for p, v in zip(ps, vals)}


def overdraw_rate(x: np.ndarray, threshold: float = 1.0) -> float:
    x = np.asarray(x, np.float64)
    return float((x > threshold).mean()) if x.size else float("nan")                                 # :291-297


def score_pairs(gt_masks: List[np.ndarray], gen_masks: List[np.ndarray], sigma: float = 2.0, device=None,
                batch: int = 4096) -> List[Dict[str, float]]:
    """compute_metrics for every pair, batched by image size on the GPU (the masks are already binary)."""
    dev = _device(device)
    out: List[Optional[Dict[str, float]]] = [None] * len(gt_masks)
    by_shape: Dict[Tuple[int, ...], List[int]] = {}
    for i, (g, p) in enumerate(zip(gt_masks, gen_masks)):
        if g.shape != p.shape:
            raise ValueError(f"Shape mismatch: gt{g.shape} vs pred{p.shape}")
        by_shape.setdefault(g.shape, []).append(i)
    for idx in by_shape.values():
        for k in range(0, len(idx), batch):
            part = idx[k:k + batch]
            g = torch.from_numpy(np.stack([gt_masks[i] for i in part]).astype(np.uint8)).to(dev)
            p = torch.from_numpy(np.stack([gen_masks[i] for i in part]).astype(np.uint8)).to(dev)
            for i, m in zip(part, compute_metrics_batch(g, p, sigma=sigma)):
                out[i] = m
    return out   # type: ignore[return-value]


def summarise(detail, missing: int, threshold: int, invert: bool, sigma: float, run_dir: Path) -> Dict[str, object]:
    """The one row of metrics_summary.csv (:413-447), in the reference's column order."""
    col = {k: np.asarray([r[k] for r in detail], np.float64) for k in ("iou", "gt_iou", "far_noise_ratio", "gauss_recall")}
    row: Dict[str, object] = {"n_pairs": len(detail), "missing_pairs_skipped": int(missing), "threshold": int(threshold),
                              "invert": bool(invert), "sigma": float(sigma), "distance_backend": DISTANCE_BACKEND}
    for k in ("iou", "gt_iou", "far_noise_ratio"):
        row[f"{k}_mean"], row[f"{k}_std"] = mean_std(col[k])
    q = quantiles(col["far_noise_ratio"], [50, 90, 95])
    row.update(far_noise_ratio_median=q["p50"], far_noise_ratio_p90=q["p90"], far_noise_ratio_p95=q["p95"])
    row["gauss_recall_mean"], row["gauss_recall_std"] = mean_std(col["gauss_recall"])
    q = quantiles(col["gauss_recall"], [50, 90, 95])
    row.update(gauss_recall_median=q["p50"], gauss_recall_p90=q["p90"], gauss_recall_p95=q["p95"],
               gauss_overdraw_rate_gt1=overdraw_rate(col["gauss_recall"], 1.0), run_dir=str(run_dir))
    return row


def main(argv=None, device=None):
    import argparse
    import pandas as pd
    ap = argparse.ArgumentParser()
    ap.add_argument("--gt_dir", type=str, required=True, help="ground-truth images (p00000.jpg ...)")
    ap.add_argument("--gen_dir", type=str, required=True, help="generated images (pic1.png ...)")
    ap.add_argument("--out_dir", type=str, required=True, help="a run_* directory is created below it")
    ap.add_argument("--threshold", type=int, default=128, help="binarisation threshold (0-255)")
    ap.add_argument("--invert", action="store_true", help="dark pixels are foreground (line drawings)")
    ap.add_argument("--sigma", type=float, default=2.0, help="sigma (px) of the Gaussian-weighted recall")
    ap.add_argument("--max_pairs", type=int, default=-1, help="score at most this many pairs (-1: all)")
    ap.add_argument("--save_diff", action="store_true", help="also save the red / blue / black difference images")
    args = ap.parse_args(argv)

    gt_dir, gen_dir, out_root = Path(args.gt_dir), Path(args.gen_dir), Path(args.out_dir)
    out_root.mkdir(parents=True, exist_ok=True)
    for what, d in (("gt_dir", gt_dir), ("gen_dir", gen_dir)):
        if not d.exists():
            raise FileNotFoundError(f"{what} not found: {d}")
    run_dir = out_root / ("run_" + datetime.now().strftime("%Y%m%d_%H%M%S"))
    sub = {k: run_dir / "binarized" / k for k in ("gt", "gen", "pair")}
    if args.save_diff:
        sub["diff"] = run_dir / "diff"
    for d in sub.values():
        d.mkdir(parents=True, exist_ok=True)

    pairs, missing = pair_files(gt_dir, gen_dir, args.max_pairs)
    if not pairs:
        raise RuntimeError("no pairs to compare: ground truth p00000.jpg, p00001.jpg ... is matched with "
                           "generated pic1.png, pic2.png ... (p00000 <-> pic1, p00001 <-> pic2, ...)")

    with ThreadPoolExecutor() as pool:     # PIL releases the GIL while it decodes
        load = lambda p: load_binary_mask(p, threshold=args.threshold, invert=args.invert)
        gt_masks = list(pool.map(load, [p[1] for p in pairs]))
        gen_masks = list(pool.map(load, [p[2] for p in pairs]))
        metrics = score_pairs(gt_masks, gen_masks, sigma=args.sigma, device=device)

        def write_artifacts(k: int) -> Dict[str, str]:
            gt_idx, gt_path, gen_path = pairs[k]
            files = {"gt_bin": sub["gt"] / f"{gt_path.stem}_bin.png", "gen_bin": sub["gen"] / f"{gen_path.stem}_bin.png",
                     "pair_bin": sub["pair"] / f"pair_gt{gt_idx:05d}_vs_{gen_path.stem}.png"}
            mask_to_pil(gt_masks[k]).save(files["gt_bin"])
            mask_to_pil(gen_masks[k]).save(files["gen_bin"])
            save_side_by_side(gt_masks[k], gen_masks[k], files["pair_bin"])
            if args.save_diff:
                files["diff_bin"] = sub["diff"] / f"diff_gt{gt_idx:05d}_vs_{gen_path.stem}.png"
                save_diff_visual(gt_masks[k], gen_masks[k], files["diff_bin"])
            rel = {c: str(f.relative_to(run_dir)) for c, f in files.items()}
            rel.setdefault("diff_bin", "")
            return rel
        written = list(pool.map(write_artifacts, range(len(pairs))))

    detail = [{"gt_index": gt_idx, "gt_file": gt_path.name, "gen_file": gen_path.name,
               **{c: written[k][c] for c in ("gt_bin", "gen_bin", "pair_bin", "diff_bin")}, **metrics[k]}
              for k, (gt_idx, gt_path, gen_path) in enumerate(pairs)]
    df = pd.DataFrame(detail)
    summary = pd.DataFrame([summarise(detail, missing, args.threshold, args.invert, args.sigma, run_dir)])
    detail_path, summary_path = run_dir / "metrics_detail.csv", run_dir / "metrics_summary.csv"
    df.to_csv(detail_path, index=False, encoding="utf-8-sig")
    summary.to_csv(summary_path, index=False, encoding="utf-8-sig")
    notes = {"gt_dir": gt_dir, "gen_dir": gen_dir, "threshold": args.threshold, "invert": bool(args.invert),
             "sigma": args.sigma, "distance_backend": DISTANCE_BACKEND, "max_pairs": args.max_pairs,
             "save_diff": bool(args.save_diff), "missing_pairs_skipped": missing}
    (run_dir / "config.txt").write_text("".join(f"{k}={v}\n" for k, v in notes.items()), encoding="utf-8")
    print(f"[OK] run_dir: {run_dir}")
    print(f"[OK] detail:  {detail_path}")
    print(f"[OK] summary: {summary_path}")
    print(summary.to_string(index=False))
    return run_dir


if __name__ == "__main__":
    main()
