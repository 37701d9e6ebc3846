"""ctypes binding of libdmb200.so -- the C-ABI library of hand-written sm_100a kernels.

PyTorch is used here for device memory, streams and (in bench.py) torch.distributed only;
all arithmetic of the sampling path runs in the library.  There is no CPU fallback: every
entry point raises if the library or a B200 is missing.
"""
from .native import (Context, NativeError, context_for, library_path, load_library,  # noqa: F401
                     PRECISION_BF16, PRECISION_TF32, write_png_batch)
