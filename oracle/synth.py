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
