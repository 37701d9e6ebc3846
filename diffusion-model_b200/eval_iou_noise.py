"""Drop-in for the scoring functions of the reference's eval_iou_noise.py, computed on the B200
(SURVEY.md 8(f) N2).  Same names, arguments, return keys and exceptions as the reference:

  load_binary_mask(image_path, threshold=128, invert=True)     eval_iou_noise.py:77-94
  gaussian_weighted_recall(gt, pred, sigma=2.0)                eval_iou_noise.py:185-208
  far_noise_ratio(gt, pred, sigma=2.0)                         eval_iou_noise.py:211-232
  compute_metrics(gt, pred, sigma=2.0) -> dict                 eval_iou_noise.py:239-272

plus compute_metrics_batch(), which scores a whole batch of decoded u8 images that is already in
HBM (what the 4096-sample config needs: the reference's per-image scipy distance transform on
the host is the wall-clock tail there).  The device returns exact integer counters and one
float64 sum per image (dmb_score_masks); the ratios are formed here in float64 exactly as the
reference forms them, so iou / gt_iou / far_noise_ratio are bit-identical and gauss_recall
differs only by the order of a float64 sum.  There is no CPU implementation in this module."""
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
