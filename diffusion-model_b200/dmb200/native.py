"""Loader and thin typed wrapper for the C ABI declared in include/dmb200.h."""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Dict, Optional, Sequence, Tuple

import torch

PRECISION_BF16 = 0
PRECISION_TF32 = 1
_PRECISION_NAMES = {"bf16": PRECISION_BF16, "tf32": PRECISION_TF32, "fp32": PRECISION_TF32}

_HERE = os.path.dirname(os.path.abspath(__file__))


class NativeError(RuntimeError):
    """A non-zero status from libdmb200 (message from dmb_last_error_string)."""


def library_path() -> str:
    return os.environ.get("DMB200_LIBRARY", os.path.join(_HERE, "libdmb200.so"))


_lib = None
_lib_lock = threading.Lock()

# name -> (restype, argtypes); mirrors include/dmb200.h one to one.
_P = C.c_void_p
_SIGNATURES = {
    "dmb_abi_version": (C.c_int, []),
    "dmb_last_error_string": (C.c_char_p, []),
    "dmb_ctx_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(_P)]),
    "dmb_ctx_destroy": (C.c_int, [_P]),
    "dmb_load_unet_weights": (C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(_P), C.POINTER(C.c_int64)]),
    "dmb_load_vae_dec_weights": (C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(_P), C.POINTER(C.c_int64)]),
    "dmb_unet_forward": (C.c_int, [_P, _P, _P, _P, _P, _P, C.c_int, _P, _P, _P]),
    "dmb_denoise_cond_step": (C.c_int, [_P, _P, _P, _P, C.c_int, _P, _P, C.c_float, _P, C.c_int, _P, C.c_uint64,
                                        C.c_uint64, _P, _P, C.c_int, _P]),
    "dmb_sample_loop": (C.c_int, [_P, _P, _P, C.c_int, _P, _P, C.c_float, _P, C.c_int, C.c_int, C.c_int, _P,
                                  C.c_uint64, C.c_uint64, _P, _P, C.c_int, _P]),
    "dmb_sample_loop_strided": (C.c_int, [_P, _P, _P, C.c_int, _P, _P, C.c_float, _P, C.c_int, C.c_int, C.c_int,
                                          C.c_int, _P, C.c_uint64, C.c_uint64, _P, _P, C.c_int, _P]),
    "dmb_vae_decode": (C.c_int, [_P, _P, C.c_int, _P, _P, _P]),
    "dmb_load_vae_enc_weights": (C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(_P), C.POINTER(C.c_int64)]),
    "dmb_vae_encode": (C.c_int, [_P, _P, _P, C.c_int, _P, _P, _P, _P, _P]),
    "dmb_add_noise": (C.c_int, [_P, _P, _P, _P, C.c_int, _P, C.c_int, C.c_int64, _P, _P]),
    "dmb_recon_error": (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P, _P, _P]),
    "dmb_kl_per_sample_static": (C.c_int, [C.c_int, _P, _P, C.c_int, C.c_int64, C.c_float, _P, _P]),
    "dmb_score_masks": (C.c_int, [_P, _P, C.c_int, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double,
                                  _P, _P, _P]),
    "dmb_write_png_batch": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.c_int, C.c_int]),
    "dmb_philox_normal": (C.c_int, [_P, _P, C.c_int, C.c_uint64, C.c_uint64, C.c_int, _P]),
    "dmb_launch_count": (C.c_int64, [_P]),
    "dmb_reset_launch_count": (C.c_int, [_P]),
    "dmb_set_kernel_timing": (C.c_int, [_P, C.c_int]),
    "dmb_get_kernel_timing": (C.c_int, [_P, C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]),
    "dmb_debug_get_tap": (C.c_int, [_P, C.c_char_p, _P, C.c_int64, C.POINTER(C.c_int64), _P]),
    "dmb_op_conv3x3": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "dmb_op_linear": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "dmb_op_conv_transpose4x4": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def load_library():
    """dlopen libdmb200.so and set the prototypes.  Raises if the library is absent:
    the product has no other implementation of the path."""
    global _lib
    with _lib_lock:
        if _lib is None:
            path = library_path()
            if not os.path.isfile(path):
                raise NativeError(
                    f"libdmb200.so not found at {path}; build it with "
                    "`python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
            lib = C.CDLL(path)
            for name, (res, args) in _SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def write_png_batch(images: torch.Tensor, paths: Sequence[str], threads: int = 0, compression_level: int = 3) -> None:
    """(n,3,H,W) / (n,1,H,W) / (n,H,W) uint8 tensor (CUDA or CPU) -> one PNG per image, encoded in
    parallel on the host by the native library (dmb_write_png_batch).  A CUDA batch crosses PCIe
    in ONE copy into pinned memory."""
    if images.dtype != torch.uint8 or images.dim() not in (3, 4):
        raise ValueError(f"images must be uint8 (n,C,H,W) or (n,H,W), got {images.dtype} {tuple(images.shape)}")
    if images.dim() == 3:
        images = images[:, None]
    n, Cc, H, W = images.shape
    if Cc not in (1, 3):
        raise ValueError(f"images must have 1 or 3 channels, got {Cc}")
    if len(paths) != n:
        raise ValueError(f"{n} images but {len(paths)} paths")
    if images.is_cuda:
        host = torch.empty(images.shape, dtype=torch.uint8, pin_memory=True)
        host.copy_(images.contiguous(), non_blocking=True)
        torch.cuda.current_stream(images.device).synchronize()
    else:
        host = images.contiguous()
    arr = (C.c_char_p * n)(*[os.fsencode(p) for p in paths])
    lib = load_library()
    rc = lib.dmb_write_png_batch(C.c_void_p(host.data_ptr()), n, Cc, H, W, arr, int(threads), int(compression_level))
    if rc != 0:
        raise NativeError(f"dmb_write_png_batch failed ({rc}): " + lib.dmb_last_error_string().decode("utf-8", "replace"))


KL_STATIC_MAX_B = 4096   # DMB_KL_STATIC_MAX_B in include/dmb200.h


def kl_per_sample_static(mu: torch.Tensor, logvar: torch.Tensor, out: torch.Tensor, kl_div: float = 224.0 * 224.0):
    """dmb_kl_per_sample_static: the per-sample KL term of VAE.encode from (B, ...) fp32 CUDA
    tensors mu / logvar into the caller's `out` (B,) -- no allocation anywhere on the path (the
    scratch is a static device array inside the library; at most KL_STATIC_MAX_B samples per call;
    calls that overlap on the GPU race).  Enqueued on torch's current stream."""
    dev = mu.device
    for name, t in (("mu", mu), ("logvar", logvar), ("out", out)):
        _check_dev(t, dev, torch.float32, name)
    if dev.type != "cuda":
        raise NativeError(f"no CPU fallback (got device '{dev}')")
    B = mu.shape[0]
    if logvar.shape != mu.shape or out.numel() != B:
        raise ValueError("mu / logvar must have one shape and out one entry per sample")
    lib = load_library()
    rc = lib.dmb_kl_per_sample_static(dev.index, _ptr(mu), _ptr(logvar), B, mu.numel() // B, float(kl_div), _ptr(out),
                                      C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    if rc != 0:
        raise NativeError(f"dmb_kl_per_sample_static failed ({rc}): " + lib.dmb_last_error_string().decode("utf-8", "replace"))
    return out


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _check_dev(t: Optional[torch.Tensor], device: torch.device, dtype, name: str):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if t.device != device:
        raise ValueError(f"{name} lives on {t.device}, context is bound to {device}")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


class Context:
    """One native context = one GPU, one precision mode.  Not thread-safe."""

    def __init__(self, device, precision: str = "bf16"):
        device = torch.device(device)
        if device.type != "cuda":
            raise NativeError("the sampling path runs on a B200 only: no CPU fallback (got device "
                              f"'{device}')")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if precision not in _PRECISION_NAMES:
            raise ValueError(f"precision must be one of {sorted(_PRECISION_NAMES)}")
        self.lib = load_library()
        self.device = device
        self.precision = precision
        self._h = C.c_void_p()
        self._call("dmb_ctx_create", device.index, _PRECISION_NAMES[precision], C.byref(self._h))
        self._unet_token = None
        self._vae_token = None
        self._vae_enc_token = None
        self._param_cache = {}

    # -- plumbing
    def _call(self, name, *args):
        rc = getattr(self.lib, name)(*args)
        if rc != 0:
            msg = self.lib.dmb_last_error_string().decode("utf-8", "replace")
            raise NativeError(f"{name} failed ({rc}): {msg}")

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.dmb_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- weights
    def _load(self, fn: str, tensors: Dict[str, torch.Tensor]):
        names = list(tensors)
        keep = [tensors[k].detach().to(torch.float32).contiguous() for k in names]
        n = len(names)
        c_names = (C.c_char_p * n)(*[k.encode() for k in names])
        c_ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in keep])
        c_numel = (C.c_int64 * n)(*[t.numel() for t in keep])
        torch.cuda.synchronize(self.device)
        self._call(fn, self._h, n, c_names, c_ptrs, c_numel)

    # The parameter walk (module.state_dict: ~0.45 ms for the U-Net's 191 tensors) is done once per
    # module "epoch"; afterwards a call checks (data_ptr, version) of the cached Parameter objects
    # (~45 us).  The drop-in modules bump `_native_epoch` in `_apply` (.to / .cuda / .half ...) and
    # `load_state_dict`; in-place writes bump the tensors' version counters; writes through `.data`
    # move neither and need `module.refresh_native()`.
    def _params(self, module: torch.nn.Module, select) -> Tuple[Tuple[str, ...], Tuple[torch.Tensor, ...]]:
        epoch = getattr(module, "_native_epoch", None)
        key = (id(module), select)
        hit = self._param_cache.get(key)
        if hit is None or epoch is None or hit[0] != epoch:
            sd = module.state_dict(keep_vars=True)
            names = tuple(k for k in sd if select is None or k.startswith(select))
            hit = (epoch, names, tuple(sd[k] for k in names))
            self._param_cache[key] = hit
        return hit[1], hit[2]

    def _ensure(self, module, select, attr: str, fn: str, extra=None):
        names, tensors = self._params(module, select)
        tok = (id(module), getattr(module, "_native_epoch", None)) + tuple((t.data_ptr(), t._version) for t in tensors)
        if tok != getattr(self, attr):
            payload = dict(zip(names, tensors))
            if extra:
                payload.update(extra())
            self._load(fn, payload)
            setattr(self, attr, tok)

    def invalidate(self, module: Optional[torch.nn.Module] = None):
        """Forget the uploaded weights of `module` (all modules if None): the next call re-uploads
        and repacks them.  Needed only after writes that bypass the parameters' version counters
        (`p.data.copy_(...)`, EMA updates through `.data`); load_state_dict / .to() are detected."""
        for name in ("_unet_token", "_vae_token", "_vae_enc_token"):
            tok = getattr(self, name)
            if tok is not None and (module is None or tok[0] == id(module)):
                setattr(self, name, None)
        for key in [k for k in self._param_cache if module is None or k[0] == id(module)]:
            del self._param_cache[key]

    def ensure_unet(self, module: torch.nn.Module):
        """Upload (repack) the U-Net's parameters if they changed since the last call."""
        # pos_encoding frequencies exactly as the reference computes them (unet_cond.py:156)
        self._ensure(module, None, "_unet_token", "dmb_load_unet_weights",
                     lambda: {"__inv_freq": 1.0 / (10000 ** (torch.arange(0, 256, 2).float() / 256))})

    def ensure_vae(self, module: torch.nn.Module):
        self._ensure(module, "dec.", "_vae_token", "dmb_load_vae_dec_weights")

    def ensure_vae_enc(self, module: torch.nn.Module):
        self._ensure(module, ("enc.", "to_mu.", "to_logvar."), "_vae_enc_token", "dmb_load_vae_enc_weights")

    # -- compute
    def unet_forward(self, x, t, y, vals, mask, want_geom=True):
        B = x.shape[0]
        x = _check_dev(x.contiguous(), self.device, torch.float32, "x")
        t = t.to(torch.long).contiguous()
        y = y.to(torch.long).contiguous()
        vals = None if vals is None else vals.to(torch.float32).contiguous()
        mask = None if mask is None else mask.to(torch.float32).contiguous()
        eps = torch.empty((B, 4, 28, 28), device=self.device, dtype=torch.float32)
        geom = torch.empty((B, 12), device=self.device, dtype=torch.float32) if want_geom else None
        self._call("dmb_unet_forward", self._h, _ptr(x), _ptr(t), _ptr(y), _ptr(vals), _ptr(mask), B, _ptr(eps),
                   _ptr(geom), self._stream())
        return eps, geom

    def denoise_cond_step(self, x, t, y, null_label, vals, mask, guidance, coef, noise=None, seed=0,
                          sample_offset=0, want_eps=False):
        B = x.shape[0]
        x = _check_dev(x.contiguous(), self.device, torch.float32, "x")
        coef = _check_dev(coef, self.device, torch.float32, "coef")
        noise = _check_dev(noise, self.device, torch.float32, "noise")
        t = t.to(torch.long).contiguous()
        y = y.to(torch.long).contiguous()
        vals = None if vals is None else vals.to(torch.float32).contiguous()
        mask = None if mask is None else mask.to(torch.float32).contiguous()
        out = torch.empty_like(x)
        eps = torch.empty_like(x) if want_eps else None
        self._call("dmb_denoise_cond_step", self._h, _ptr(x), _ptr(t), _ptr(y), int(null_label), _ptr(vals),
                   _ptr(mask), float(guidance), _ptr(coef), coef.shape[0], _ptr(noise), int(seed),
                   int(sample_offset), _ptr(out), _ptr(eps), B, self._stream())
        return (out, eps) if want_eps else out

    def sample_loop(self, x, y, null_label, vals, mask, guidance, coef, t_start, n_steps, noise=None, seed=0,
                    sample_offset=0, x_dump=None, eps_dump=None, t_stride=1):
        """In place on x (B,4,28,28).  t_stride > 1: iteration k runs at t = t_start - k * t_stride."""
        B = x.shape[0]
        _check_dev(x, self.device, torch.float32, "x")
        coef = _check_dev(coef, self.device, torch.float32, "coef")
        for nm, tt in (("noise", noise), ("x_dump", x_dump), ("eps_dump", eps_dump)):
            _check_dev(tt, self.device, torch.float32, nm)
            if tt is not None and tuple(tt.shape) != (n_steps, B, 4, 28, 28):
                raise ValueError(f"{nm} must have shape ({n_steps},{B},4,28,28), got {tuple(tt.shape)}")
        y = y.to(torch.long).contiguous()
        vals = None if vals is None else vals.to(torch.float32).contiguous()
        mask = None if mask is None else mask.to(torch.float32).contiguous()
        if int(t_stride) == 1:
            self._call("dmb_sample_loop", self._h, _ptr(x), _ptr(y), int(null_label), _ptr(vals), _ptr(mask),
                       float(guidance), _ptr(coef), coef.shape[0], int(t_start), int(n_steps), _ptr(noise), int(seed),
                       int(sample_offset), _ptr(x_dump), _ptr(eps_dump), B, self._stream())
        else:
            self._call("dmb_sample_loop_strided", self._h, _ptr(x), _ptr(y), int(null_label), _ptr(vals), _ptr(mask),
                       float(guidance), _ptr(coef), coef.shape[0], int(t_start), int(t_stride), int(n_steps),
                       _ptr(noise), int(seed), int(sample_offset), _ptr(x_dump), _ptr(eps_dump), B, self._stream())
        return x

    def vae_decode(self, z, want_f32=True, want_u8=False):
        B = z.shape[0]
        z = _check_dev(z.to(torch.float32).contiguous(), self.device, torch.float32, "z")
        f32 = torch.empty((B, 3, 224, 224), device=self.device, dtype=torch.float32) if want_f32 else None
        u8 = torch.empty((B, 3, 224, 224), device=self.device, dtype=torch.uint8) if want_u8 else None
        self._call("dmb_vae_decode", self._h, _ptr(z), B, _ptr(f32), _ptr(u8), self._stream())
        return f32, u8

    def vae_encode(self, img, eps, want_moments=False):
        """VAE.encode on the device (vae.py:51-62).  img (B,3,224,224) fp32, eps (B,4,28,28) fp32 CUDA
        tensors.  Returns (z, kl_per_sample, mu, logvar); mu / logvar are None unless want_moments."""
        B = img.shape[0]
        img = _check_dev(img.to(torch.float32).contiguous(), self.device, torch.float32, "x")
        eps = _check_dev(eps.to(torch.float32).contiguous(), self.device, torch.float32, "eps")
        if tuple(eps.shape) != (B, 4, 28, 28):
            raise ValueError(f"eps must be ({B},4,28,28), got {tuple(eps.shape)}")
        z = torch.empty((B, 4, 28, 28), device=self.device, dtype=torch.float32)
        mu = torch.empty_like(z) if want_moments else None
        logvar = torch.empty_like(z) if want_moments else None
        kl = torch.empty((B,), device=self.device, dtype=torch.float32)
        self._call("dmb_vae_encode", self._h, _ptr(img), _ptr(eps), B, _ptr(z), _ptr(mu), _ptr(logvar), _ptr(kl),
                   self._stream())
        return z, kl, mu, logvar

    def add_noise(self, x0, t, alpha_bars, noise):
        """Diffuser.add_noise on the device (diff.py:18-30); bit-identical for the same noise."""
        B = x0.shape[0]
        x0 = _check_dev(x0.contiguous(), self.device, torch.float32, "x_0")
        noise = _check_dev(noise.contiguous(), self.device, torch.float32, "noise")
        if noise.shape != x0.shape:
            raise ValueError(f"noise {tuple(noise.shape)} does not match x_0 {tuple(x0.shape)}")
        t = _check_dev(t.to(torch.long).contiguous(), self.device, torch.long, "t")
        if t.numel() != B:
            raise ValueError(f"t must hold one timestep per sample ({B}), got {t.numel()}")
        ab = _check_dev(alpha_bars.to(torch.float32).contiguous(), self.device, torch.float32, "alpha_bars")
        out = torch.empty_like(x0)
        self._call("dmb_add_noise", self._h, _ptr(x0), _ptr(t), _ptr(ab), int(ab.numel()), _ptr(noise), B,
                   int(x0.numel() // B), _ptr(out), self._stream())
        return out

    def recon_error(self, a, b):
        """Per-sample float64 sums of (a - b)^2 and |a - b| (dmb_recon_error): the reconstruction term of
        VAE.forward (vae.py:71-76) and the per-image errors of vae_recon_check.py:29-31."""
        a = _check_dev(a.contiguous(), self.device, torch.float32, "a")
        b = _check_dev(b.contiguous(), self.device, torch.float32, "b")
        if a.shape != b.shape:
            raise ValueError(f"shapes differ: {tuple(a.shape)} vs {tuple(b.shape)}")
        B = a.shape[0]
        sq = torch.empty((B,), device=self.device, dtype=torch.float64)
        ab = torch.empty((B,), device=self.device, dtype=torch.float64)
        self._call("dmb_recon_error", self._h, _ptr(a), _ptr(b), B, int(a.numel() // B), _ptr(sq), _ptr(ab),
                   self._stream())
        return sq, ab

    def score_masks(self, pred, gt, threshold=128, invert=True, sigma=2.0):
        """Device-side counters of eval_iou_noise.compute_metrics for a batch.  pred / gt: uint8 CUDA
        tensors, (B,3,H,W) RGB images or (B,1,H,W) grey images (binarised like load_binary_mask) or
        (B,H,W) masks (non-zero = foreground).  Returns (counts (B,6) int64, weighted_hit (B,) f64)."""
        def chan(t, name):
            _check_dev(t, self.device, torch.uint8, name)
            if t.dim() == 3:
                return 0
            if t.dim() == 4 and t.shape[1] in (1, 3):
                return int(t.shape[1])
            raise ValueError(f"{name} must be (B,H,W), (B,1,H,W) or (B,3,H,W) uint8, got {tuple(t.shape)}")
        pc, gc = chan(pred, "pred"), chan(gt, "gt")
        pred, gt = pred.contiguous(), gt.contiguous()
        B, H, W = pred.shape[0], pred.shape[-2], pred.shape[-1]
        if (gt.shape[0], gt.shape[-2], gt.shape[-1]) != (B, H, W):
            raise ValueError(f"Shape mismatch: gt{tuple(gt.shape)} vs pred{tuple(pred.shape)}")
        counts = torch.empty((B, 6), device=self.device, dtype=torch.int64)
        hit = torch.empty((B,), device=self.device, dtype=torch.float64)
        self._call("dmb_score_masks", self._h, _ptr(pred), pc, _ptr(gt), gc, B, H, W, int(threshold), int(bool(invert)),
                   float(sigma), _ptr(counts), _ptr(hit), self._stream())
        return counts, hit

    def philox_normal(self, B, seed, sample_offset, t):
        out = torch.empty((B, 4, 28, 28), device=self.device, dtype=torch.float32)
        self._call("dmb_philox_normal", self._h, _ptr(out), B, int(seed), int(sample_offset), int(t), self._stream())
        return out

    def launch_count(self) -> int:
        return int(self.lib.dmb_launch_count(self._h))

    def reset_launch_count(self):
        self._call("dmb_reset_launch_count", self._h)

    def set_kernel_timing(self, on: bool):
        self._call("dmb_set_kernel_timing", self._h, 1 if on else 0)

    def kernel_timing(self):
        """[{label, count, ms, flops, bytes}] of the eager calls made since set_kernel_timing(True)."""
        import json
        need = C.c_int64(0)
        self._call("dmb_get_kernel_timing", self._h, None, 0, C.byref(need))
        buf = C.create_string_buffer(need.value)
        self._call("dmb_get_kernel_timing", self._h, buf, need.value, C.byref(need))
        return json.loads(buf.value.decode())

    def get_tap(self, name: str, capacity_elems: int = 1 << 26):
        buf = torch.empty((capacity_elems,), device=self.device, dtype=torch.float32)
        shape = (C.c_int64 * 4)()
        self._call("dmb_debug_get_tap", self._h, name.encode(), _ptr(buf), capacity_elems, shape, self._stream())
        n, c, h, w = (int(v) for v in shape)
        return buf[: n * c * h * w].view(n, c, h, w).clone()

    def op_conv3x3(self, x, w):
        B, Cin, H, W = x.shape
        Cout = w.shape[0]
        x = _check_dev(x.contiguous(), self.device, torch.float32, "x")
        w = w.to(torch.float32).contiguous()
        out = torch.empty((B, Cout, H, W), device=self.device, dtype=torch.float32)
        self._call("dmb_op_conv3x3", self._h, _ptr(x), _ptr(w), B, Cin, Cout, H, W, _ptr(out), self._stream())
        return out

    def op_linear(self, x, w, bias=None, act=0):
        rows, K = x.shape
        N = w.shape[0]
        x = x.to(torch.float32).contiguous()
        w = w.to(torch.float32).contiguous()
        bias = None if bias is None else bias.to(torch.float32).contiguous()
        out = torch.empty((rows, N), device=self.device, dtype=torch.float32)
        self._call("dmb_op_linear", self._h, _ptr(x), _ptr(w), _ptr(bias), rows, K, N, int(act), _ptr(out),
                   self._stream())
        return out

    def op_conv_transpose4x4(self, x, w, bias=None):
        B, Cin, H, W = x.shape
        Cout = w.shape[1]
        x = _check_dev(x.contiguous(), self.device, torch.float32, "x")
        w = w.to(torch.float32).contiguous()
        bias = None if bias is None else bias.to(torch.float32).contiguous()
        out = torch.empty((B, Cout, 2 * H, 2 * W), device=self.device, dtype=torch.float32)
        self._call("dmb_op_conv_transpose4x4", self._h, _ptr(x), _ptr(w), _ptr(bias), B, Cin, Cout, H, W,
                   _ptr(out), self._stream())
        return out


_contexts: Dict[Tuple[int, str], Context] = {}


def invalidate_module(module: Optional[torch.nn.Module] = None) -> None:
    """Context.invalidate on every live context of this process."""
    for ctx in _contexts.values():
        ctx.invalidate(module)


def default_precision() -> str:
    return os.environ.get("DMB200_PRECISION", "bf16")


def context_for(device, precision: Optional[str] = None) -> Context:
    """The process-wide context of a device (created on first use)."""
    device = torch.device(device)
    if device.type != "cuda":
        raise NativeError(f"the sampling path runs on a B200 only: no CPU fallback (got device '{device}')")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    precision = precision or default_precision()
    key = (idx, precision)
    ctx = _contexts.get(key)
    if ctx is None:
        ctx = Context(torch.device("cuda", idx), precision)
        _contexts[key] = ctx
    return ctx
