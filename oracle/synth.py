"""Synthetic inputs shared by tests, smoke() and bench.py (SURVEY.md §8d recipe).

TEST / BENCH INFRASTRUCTURE.  No product code imports this.
"""
from __future__ import annotations

import numpy as np
import torch


def synth_rows(n: int, seed: int = 0) -> np.ndarray:
    """n rows of the headerless 13-column entity CSV (entityCsvSampler.py:14-18):
    col0 dummy; 1-4 line x1,y1,x2,y2; 5-7 circle cx,cy,r; 8-12 arc ax,ay,r,th1,th2.
    Drawing coordinates for a 400x400 sheet."""
    rng = np.random.default_rng(seed)
    rows = np.zeros((n, 13), np.float64)
    rows[:, 0] = np.arange(n)
    rows[:, 1:5] = rng.uniform(20, 380, (n, 4))
    rows[:, 5:7] = rng.uniform(100, 300, (n, 2))
    rows[:, 7] = rng.uniform(20, 90, n)
    rows[:, 8:10] = rng.uniform(100, 300, (n, 2))
    rows[:, 10] = rng.uniform(20, 90, n)
    rows[:, 11:13] = rng.uniform(0, 360, (n, 2))
    return np.round(rows, 3)


def write_csv(path: str, rows: np.ndarray) -> None:
    np.savetxt(path, rows, delimiter=",", fmt="%.3f")


def mixed_conditions(batch: int, seed: int = 0):
    """Config 3/4: line, circle and arc rows concatenated in class order
    (diff.py:223-226 orders samples by class).  Returns (y, vals, mask, counts)."""
    from .ref_torch import build_vals_mask
    n1 = (batch + 2) // 3
    n2 = (batch - n1 + 1) // 2
    n3 = batch - n1 - n2
    rows = synth_rows(batch, seed)
    parts, ys, o = [], [], 0
    for cls, n in ((1, n1), (2, n2), (3, n3)):
        if n:
            parts.append(build_vals_mask(rows[o:o + n], cls))
            ys += [cls] * n
            o += n
    vals = np.concatenate([p[0] for p in parts])
    mask = np.concatenate([p[1] for p in parts])
    counts = [(c, n) for c, n in ((1, n1), (2, n2), (3, n3)) if n]
    return (torch.tensor(ys, dtype=torch.long), torch.from_numpy(vals), torch.from_numpy(mask), counts)


def line_conditions(batch: int, seed: int = 0):
    from .ref_torch import build_vals_mask
    vals, mask = build_vals_mask(synth_rows(batch, seed), 1)
    return (torch.full((batch,), 1, dtype=torch.long), torch.from_numpy(vals), torch.from_numpy(mask))


def noise_bank(T: int, shape, seed: int = 1234):
    """x_T and the T per-step noise tensors from a dedicated CPU generator."""
    g = torch.Generator().manual_seed(seed)
    x_T = torch.randn(shape, generator=g)
    noise = torch.randn((T,) + tuple(shape), generator=g)
    return x_T, noise


def render_gt(y, vals, size: int = 224, half_width: float = 1.0, chunk: int = 64) -> np.ndarray:
    """Pseudo ground truth for the scoring configs: the entity of each condition row drawn as a dark
    stroke (0) on a white (255) `size` x `size` grey sheet, (N, size, size) uint8 -- what
    eval_iou_noise.load_binary_mask(invert=True) turns into a thin foreground mask.  Conditions are
    the normalised (N,12) vectors of entityCsvSampler.py:101-163 (x/W, 1 - y/H, r/W, angle/360), so
    pixel x = v_x * size and pixel row = v_y * size (the Y flip of the condition puts the sheet's
    origin at the bottom; an image row counts from the top, which is the same flip again).
    A pixel is stroke if its centre lies within `half_width` of the entity.  Deterministic; no RNG."""
    y = np.asarray(y)
    vals = np.asarray(vals, dtype=np.float64)
    n = y.shape[0]
    out = np.full((n, size, size), 255, np.uint8)
    py, px = np.meshgrid(np.arange(size) + 0.5, np.arange(size) + 0.5, indexing="ij")
    for b0 in range(0, n, chunk):
        yy = y[b0:b0 + chunk]
        v = vals[b0:b0 + chunk] * size
        d = np.full((len(yy), size, size), 1e9)
        X, Y = px[None], py[None]
        ln = yy == 1
        if ln.any():
            x1, y1, x2, y2 = (v[ln, i][:, None, None] for i in range(4))
            dx, dy = x2 - x1, y2 - y1
            L2 = np.maximum(dx * dx + dy * dy, 1e-12)
            u = np.clip(((X - x1) * dx + (Y - y1) * dy) / L2, 0.0, 1.0)
            d[ln] = np.hypot(X - (x1 + u * dx), Y - (y1 + u * dy))
        ci = yy == 2
        if ci.any():
            cx, cy, r = (v[ci, i][:, None, None] for i in (4, 5, 6))
            d[ci] = np.abs(np.hypot(X - cx, Y - cy) - r)
        ar = yy == 3
        if ar.any():
            ax, ay, r = (v[ar, i][:, None, None] for i in (7, 8, 9))
            th1 = vals[b0:b0 + chunk][ar, 10][:, None, None] * 360.0
            th2 = vals[b0:b0 + chunk][ar, 11][:, None, None] * 360.0
            ang = np.degrees(np.arctan2(-(Y - ay), X - ax)) % 360.0      # counter-clockwise, sheet orientation
            span = (th2 - th1) % 360.0
            inside = ((ang - th1) % 360.0) <= span
            d[ar] = np.where(inside, np.abs(np.hypot(X - ax, Y - ay) - r), 1e9)
        out[b0:b0 + chunk][d <= half_width] = 0
    return out
