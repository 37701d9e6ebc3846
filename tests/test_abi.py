"""The C-ABI library loads and exports every symbol include/dmb200.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "dmb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dmb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_header_symbols():
    from dmb200 import load_library, native
    lib = load_library()
    names = _declared()
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(native.EXPORTED_SYMBOLS) == names
    assert lib.dmb_abi_version() == 1


def test_signatures_have_no_torch_types():
    src = open(os.path.join(ROOT, "include", "dmb200.h")).read()
    assert 'extern "C"' in src
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)     # comments may cite torch calls of the reference
    for banned in ("torch", "at::", "c10::", "Tensor", "std::", "#include <cuda", "cudaStream_t"):
        assert banned not in code, banned
    includes = re.findall(r"#include\s*[<\"]([^>\"]+)[>\"]", code)
    assert includes == ["stdint.h"], includes


def test_ctx_create_without_gpu_fails_cleanly():
    import torch
    if torch.cuda.is_available():
        return
    from dmb200 import load_library
    lib = load_library()
    h = ctypes.c_void_p()
    rc = lib.dmb_ctx_create(0, 0, ctypes.byref(h))
    assert rc != 0 and h.value is None
    assert b"no CUDA device" in lib.dmb_last_error_string() or b"CUDA" in lib.dmb_last_error_string()
