"""Utils -- the checkpoint / image I/O subset of the reference's ``utils.py`` that the
sampling scripts use (utils.py:62-73, 216-224).  Training bookkeeping (loss plots,
record.txt) is outside the sampling path and is not provided; matplotlib is not imported."""
from __future__ import annotations

import os

import torch
from torch import nn


class Utils:
    @staticmethod
    def saveModelParameter(dir_path: str, model: nn.Module) -> None:
        torch.save(model.state_dict(), os.path.join(dir_path, "trained_para.pth"))

    @staticmethod
    def loadModel(path: str, model: nn.Module, device="cpu") -> nn.Module:
        """load_state_dict(torch.load(path)) -> .to(device) -> .eval()  (utils.py:68-73).
        The native weight copy is rebuilt lazily at the next forward."""
        model.load_state_dict(torch.load(path, map_location=device))
        model.to(device=device)
        model.eval()
        return model

    @staticmethod
    def saveImages(dir_path: str, images) -> None:
        """pic1.png, pic2.png, ... (utils.py:216-224)."""
        for i, image in enumerate(images):
            image.save(os.path.join(dir_path, f"pic{i + 1}.png"))
