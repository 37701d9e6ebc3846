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
    def saveImages(dir_path: str, images, start: int = 0) -> None:
        """pic1.png, pic2.png, ... (utils.py:216-224).  A list of PIL images is saved one by one
        like the reference does.  A uint8 tensor (n,3,H,W) -- what
        ``Diffuser.sample_latent_cond(..., return_u8=True)`` returns, still in HBM -- takes the
        batched path: one device->host copy into pinned memory and a parallel PNG encode in the
        native library (dmb_write_png_batch); the files decode to the same pixels.  `start`
        offsets the numbering (pic{start+i+1}.png) for callers that save in chunks."""
        if isinstance(images, torch.Tensor):
            from dmb200.native import write_png_batch
            write_png_batch(images, [os.path.join(dir_path, f"pic{start + i + 1}.png") for i in range(images.shape[0])])
            return
        for i, image in enumerate(images):
            image.save(os.path.join(dir_path, f"pic{start + i + 1}.png"))

    @staticmethod
    def saveStepImages(pixel_dir: str, images_u8: torch.Tensor, steps) -> None:
        """generate_steps.py:164-167 for a whole dump at once: ``pixel/t{i}.png`` for the decoded
        u8 images (n,3,H,W) of the timesteps in `steps`."""
        from dmb200.native import write_png_batch
        write_png_batch(images_u8, [os.path.join(pixel_dir, f"t{int(i)}.png") for i in steps])
