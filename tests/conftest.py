import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRODUCT = os.path.join(ROOT, "diffusion-model_b200")
for p in (ROOT, PRODUCT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# The tests run the PRODUCT switch set: no experiment / debug environment variable is set here.
# The one test that needs the un-normalised stage outputs ("down1" .. "up3" taps, which the
# product never stores) builds a private context under DMB200_KEEP_TAPS=1 itself
# (tests/test_gpu_model.py::test_forward_taps_and_outputs).
for _k in [k for k in os.environ if k.startswith("DMB200_") and k not in ("DMB200_LIBRARY",)]:
    del os.environ[_k]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(GOLDEN, "golden.npz"))


@pytest.fixture(scope="session")
def seeded_models():
    """Random-init product modules under the benchmark recipe:
    torch.manual_seed(0); UnetCondWithGeomHead(); VAE()."""
    import torch
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    torch.manual_seed(0)
    unet = UnetCondWithGeomHead().eval()
    vae = VAE().eval()
    return unet, vae
