"""EntityCsvSampler -- drop-in for the reference's ``entityCsvSampler.py`` (:20-199).

Reads the headerless 13-column entity CSV, builds the (N,12) ``cond_vals`` / ``cond_mask``
pair for one class and hands it to ``Diffuser.sample_latent_cond``.  This front-end is tiny,
runs once per call and stays host-side numpy; the sampling it starts
runs in libdmb200.

CSV columns: 0 dummy | 1-4 line x1,y1,x2,y2 | 5-7 circle cx,cy,r | 8-12 arc ax,ay,r,th1,th2.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import pandas as pd
import torch

from diff import Diffuser


class EntityCsvSampler:
    KEY_ORDER = ["x1", "y1", "x2", "y2", "cx", "cy", "cr", "ax", "ay", "ar", "theta1", "theta2"]
    IDX: Dict[str, int] = {k: i for i, k in enumerate(KEY_ORDER)}

    # class id -> [(key, csv column, kind)]; kind: x -> v/W, y -> 1 - v/H (Y flip), r -> v/W, a -> angle
    _LAYOUT = {
        1: [("x1", 1, "x"), ("y1", 2, "y"), ("x2", 3, "x"), ("y2", 4, "y")],
        2: [("cx", 5, "x"), ("cy", 6, "y"), ("cr", 7, "r")],
        3: [("ax", 8, "x"), ("ay", 9, "y"), ("ar", 10, "r"), ("theta1", 11, "a"), ("theta2", 12, "a")],
    }

    def __init__(self, diffuser: Diffuser, model, vae, class_id: int = 1,
                 base_wh: Optional[Tuple[float, float]] = (400, 400), device: Optional[torch.device] = None):
        self.diffuser = diffuser
        self.model = model
        self.vae = vae
        self.class_id = int(class_id)
        self.base_wh = base_wh
        self.device = device or getattr(diffuser, "device",
                                        torch.device("cuda" if torch.cuda.is_available() else "cpu"))

    def set_class_id(self, class_id: int) -> None:
        self.class_id = int(class_id)

    # ------------------------------------------------------------------ public API
    def load_cond(self, csv_path: str, count: Optional[int] = None, start: int = 0):
        """(cond_vals, cond_mask) tensors for rows [start, start+count) without sampling
        (entityCsvSampler.py:82-98)."""
        df = pd.read_csv(csv_path, header=None)
        vals_np, mask_np = self._build_vals_mask_for(df, self.class_id, self.base_wh)
        end = len(vals_np) if count is None else min(start + count, len(vals_np))
        if start >= end:
            raise ValueError("選択範囲にデータがありません。（start/countを確認）")
        vals = torch.from_numpy(vals_np[start:end]).float().to(self.device)
        mask = torch.from_numpy(mask_np[start:end]).float().to(self.device)
        return vals, mask

    def sample(self, csv_path: str, count: Optional[int] = None, start: int = 0, guidance_scale: float = 3.0):
        """CSV rows -> condition -> CFG sampling -> list of PIL images (entityCsvSampler.py:50-80)."""
        vals, mask = self.load_cond(csv_path, count=count, start=start)
        return self.diffuser.sample_latent_cond(
            model=self.model,
            class_counts=(self.class_id, vals.shape[0]),
            vae=self.vae,
            guidance_scale=guidance_scale,
            cond=vals,
            cond_mask=mask,
        )

    # ------------------------------------------------------------------ condition building
    def _build_vals_mask_for(self, df: pd.DataFrame, class_id: int, base_wh):
        """Normalised values and mask, float32 (entityCsvSampler.py:101-163)."""
        if class_id not in self._LAYOUT:
            raise ValueError("class_id must be 1(line), 2(circle), or 3(arc).")
        drawW, drawH = base_wh if base_wh is not None else self._infer_base_wh(df, class_id)
        W, H = np.float32(drawW), np.float32(drawH)
        n = len(df)
        vals = np.zeros((n, len(self.KEY_ORDER)), dtype=np.float32)
        mask = np.zeros_like(vals)
        for key, col, kind in self._LAYOUT[class_id]:
            v = df[col].to_numpy(dtype=np.float32)
            if kind == "x" or kind == "r":
                out = v / W
            elif kind == "y":
                out = 1.0 - (v / H)
            else:
                out = self._norm_angle_vec(v)
            vals[:, self.IDX[key]] = out
            mask[:, self.IDX[key]] = 1.0
        return vals, mask

    @staticmethod
    def _snap(v: float, choices=(224, 256, 280, 300, 320, 384, 400, 448), tol=1.5) -> float:
        for c in choices:
            if abs(v - c) <= tol:
                return float(c)
        return float(v)

    def _infer_base_wh(self, df: pd.DataFrame, class_id: int):
        """Sheet size guessed from the largest coordinate of the class's position columns,
        snapped to common sizes (entityCsvSampler.py:172-189)."""
        cols = {1: ([1, 3], [2, 4]), 2: ([5], [6]), 3: ([8], [9])}
        if class_id not in cols:
            raise ValueError("class_id must be 1(line), 2(circle), or 3(arc).")
        xs, ys = cols[class_id]
        x_max = float(np.max(np.abs(df[xs].to_numpy())))
        y_max = float(np.max(np.abs(df[ys].to_numpy())))
        return self._snap(x_max), self._snap(y_max)

    @staticmethod
    def _norm_angle_vec(v: np.ndarray) -> np.ndarray:
        """|v| > 1 is taken as degrees -> (v mod 360)/360; values already in [0,1] pass through
        (entityCsvSampler.py:192-199)."""
        out = v.astype(np.float32).copy()
        deg = np.abs(out) > 1.0
        out[deg] = (out[deg] % 360.0) / 360.0
        return out
