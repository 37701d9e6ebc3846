"""Diffuser -- B200-native drop-in for the reference's ``diff.py`` (diff.py:10-369).

Same class, method names, argument meaning and error behaviour for the conditional
reverse-diffusion path; the arithmetic (two U-Net passes as one doubled batch, CFG combine,
DDPM posterior update, noise injection, the whole T-step loop as CUDA-graph replays and the
VAE decode) runs in libdmb200.  The scripts ``generate_cond.py`` / ``generate_steps.py`` of
the reference import this module unchanged.  There is no CPU path.

Extensions (keyword-only, default to the reference behaviour): ``noise`` / ``seed`` on
``sample_latent_cond`` and ``noise`` on ``denoise_cond`` make runs reproducible and let the
parity tests inject fixed noise; ``latent_dump`` returns x_t before every step.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Tuple, Union

import torch

try:  # progress bars are cosmetic
    from tqdm import tqdm
except Exception:  # pragma: no cover
    tqdm = None

KEY_ORDER = ["x1", "y1", "x2", "y2", "cx", "cy", "cr", "ax", "ay", "ar", "theta1", "theta2"]
CLASS_KEYS = {1: ["x1", "y1", "x2", "y2"], 2: ["cx", "cy", "cr"], 3: ["ax", "ay", "ar", "theta1", "theta2"]}


def _is_native(model) -> bool:
    return hasattr(model, "_native")


class Diffuser:
    def __init__(self, num_timesteps=1000, beta_start=0.0001, beta_end=0.02, device="cpu"):
        # Same torch expressions as the reference (diff.py:11-16): the schedule is bit-identical.
        self.num_timesteps = num_timesteps
        self.device = device
        self.betas = torch.linspace(beta_start, beta_end, num_timesteps, device=device)
        self.alphas = 1 - self.betas
        self.alpha_bars = torch.cumprod(self.alphas, dim=0)
        self._coef = None

    # ------------------------------------------------------------------ schedule table
    def step_coefficients(self) -> torch.Tensor:
        """(T,3) fp32 per-t constants of the posterior step, formed with the reference's own
        expressions (diff.py:141-144, 160-161): (1-a)/sqrt(1-abar), sqrt(a),
        sqrt((1-a)(1-abar_prev)/(1-abar)) with abar_prev clamped at t = 1."""
        if self._coef is None:
            a, ab = self.alphas, self.alpha_bars
            idx = torch.clamp(torch.arange(self.num_timesteps, device=a.device) - 1, min=0)
            ab_prev = ab[idx]
            c1 = (1 - a) / torch.sqrt(1 - ab)
            sq = torch.sqrt(a)
            sigma = torch.sqrt((1 - a) * (1 - ab_prev) / (1 - ab))
            self._coef = torch.stack([c1, sq, sigma], dim=1).contiguous().float()
        return self._coef

    def strided_coefficients(self, sample_steps: int):
        """EXTENSION (not in the reference): constants of a deterministic
        DDIM (eta = 0) pass over the sub-sampled schedule t_k = T - k * (T // sample_steps).

        The DDIM update  x' = sqrt(ab') * (x - sqrt(1-ab) eps) / sqrt(ab) + sqrt(1-ab') eps  is
        x' = a x + b eps with a = sqrt(ab'/ab), b = sqrt(1-ab') - a sqrt(1-ab); written in the form the
        step kernel evaluates, x' = (x - c0 eps) / c1 + c2 z, that is c0 = -b/a, c1 = 1/a, c2 = 0
        (ab' = alpha_bar of the NEXT visited timestep, 1 after the last).  Returns (coef (T,3), stride)."""
        T = self.num_timesteps
        if sample_steps < 1 or T % sample_steps != 0:
            raise ValueError(f"sample_steps must divide num_timesteps={T}, got {sample_steps}")
        stride = T // sample_steps
        ab = self.alpha_bars.double()
        t = torch.arange(T, 0, -stride, device=ab.device)            # visited timesteps T, T-s, ..., s
        t_prev = t - stride                                           # >= 0
        ab_t = ab[t - 1]
        ab_p = torch.where(t_prev >= 1, ab[torch.clamp(t_prev - 1, min=0)], torch.ones_like(ab_t))
        a = torch.sqrt(ab_p / ab_t)
        b = torch.sqrt(1 - ab_p) - a * torch.sqrt(1 - ab_t)
        coef = torch.zeros((T, 3), dtype=torch.float64, device=ab.device)
        coef[:, 1] = 1.0
        coef[t - 1, 0] = -b / a
        coef[t - 1, 1] = 1.0 / a
        return coef.float().contiguous(), stride

    def _ctx(self, model):
        from dmb200 import context_for
        dev = torch.device(self.device)
        ctx = context_for(dev, getattr(model, "precision", None))
        ctx.ensure_unet(model)
        return ctx

    # ------------------------------------------------------------------ image conversion
    def reverse_to_img(self, x):
        """x*255 -> clamp -> uint8 (truncation) -> host -> PIL  (diff.py:58-64)."""
        from PIL import Image
        x = x * 255
        x = x.clamp(0, 255)
        x = x.to(torch.uint8).cpu()
        return _chw_u8_to_pil(x)

    # ------------------------------------------------------------------ one reverse step
    def denoise_cond(self, model, x, t, y=None, guidance_scale=0.0, null_label=0, cond_vals=None,
                     cond_mask=None, *, noise=None):
        """One DDPM step with classifier-free guidance (diff.py:127-162)."""
        T = self.num_timesteps
        # the reference's range assertion (diff.py:140) and nn.Embedding's index check, one sync
        ok = [((t >= 1) & (t <= T)).all()]
        if y is not None:
            ok.append(((y >= 0) & (y <= 3)).all().to(ok[0].device))
        ok = torch.stack(ok).tolist()
        assert ok[0]
        if not all(ok):
            raise IndexError("index out of range in self")
        if not (guidance_scale and y is not None and guidance_scale > 0):
            # The reference has no working non-CFG branch here: with y given it never assigns
            # eps (diff.py:152-156), with y=None it multiplies a tensor by the model's tuple.
            if y is not None:
                raise UnboundLocalError("cannot access local variable 'eps' where it is not associated with a "
                                        "value (reference behaviour of denoise_cond with guidance_scale <= 0)")
            raise TypeError("denoise_cond without guidance is not supported by UnetCondWithGeomHead, whose "
                            "forward returns a tuple (reference behaviour)")
        if not _is_native(model):
            raise TypeError("denoise_cond needs the B200-native UnetCondWithGeomHead / UnetCond from this package")
        ctx = self._ctx(model)
        if (cond_vals is None) or (cond_mask is None):
            cond_vals = cond_mask = None
        if noise is None:
            # same draw as the reference: torch's generator of the tensor's device (diff.py:158)
            noise = torch.randn_like(x, device=self.device)
        coef = self.step_coefficients()
        return ctx.denoise_cond_step(x.float(), t, y, null_label, cond_vals, cond_mask, guidance_scale, coef,
                                     noise=noise.float().contiguous())

    # ------------------------------------------------------------------ loops
    def _run_loop(self, ctx, x, y, vals, msk, guidance_scale, null_label, progress, noise, seed, sample_offset,
                  latent_dump, desc, sample_steps=None):
        if sample_steps is not None:
            # extension: strided deterministic schedule (see strided_coefficients)
            coef, stride = self.strided_coefficients(int(sample_steps))
            if noise is not None:
                raise ValueError("the strided (eta = 0) schedule draws no noise")
            B = x.shape[0]
            dump = None
            if latent_dump:
                dump = torch.empty((int(sample_steps),) + tuple(x.shape), device=x.device, dtype=torch.float32)
            ctx.sample_loop(x, y, null_label, vals, msk, guidance_scale, coef, self.num_timesteps, int(sample_steps),
                            seed=seed, sample_offset=sample_offset, x_dump=dump, t_stride=stride)
            return x, dump
        T = self.num_timesteps
        coef = self.step_coefficients()
        B = x.shape[0]
        if noise is not None and tuple(noise.shape) != (T, B) + tuple(x.shape[1:]):
            raise ValueError(f"noise must have shape {(T, B) + tuple(x.shape[1:])}, got {tuple(noise.shape)}")
        dump = None
        if latent_dump:
            dump = torch.empty((T,) + tuple(x.shape), device=x.device, dtype=torch.float32)
        # The bar advances per enqueued chunk of steps; nothing synchronises per step.
        chunk = 50 if progress else T
        bar = tqdm(total=T, desc=desc) if (progress and tqdm is not None) else None
        done = 0
        while done < T:
            n = min(chunk, T - done)
            ctx.sample_loop(x, y, null_label, vals, msk, guidance_scale, coef, T - done, n,
                            noise=None if noise is None else noise[done:done + n],
                            seed=seed, sample_offset=sample_offset,
                            x_dump=None if dump is None else dump[done:done + n])
            done += n
            if bar is not None:
                bar.update(n)
        if bar is not None:
            bar.close()
        return x, dump

    def sample_cond(self, model, x_shape, y, guidance_scale=0.0, null_label=0):
        """Latent-only CFG loop (diff.py:165-172)."""
        batch_size = x_shape[0]
        assert y.shape[0] == batch_size
        if not (guidance_scale and guidance_scale > 0):
            raise UnboundLocalError("cannot access local variable 'eps' where it is not associated with a value "
                                    "(reference behaviour of denoise_cond with guidance_scale <= 0)")
        ctx = self._ctx(model)
        x = torch.randn(x_shape, device=self.device)
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        x, _ = self._run_loop(ctx, x, y.to(self.device), None, None, guidance_scale, null_label, False, None, seed,
                              0, False, "")
        return x

    def sample_latent_cond(
        self,
        model,
        class_counts: Union[Dict[int, int], Tuple[int, int], List[Tuple[int, int]]],
        z_shape: Tuple[int, int, int] = None,
        vae=None,
        to_pil: bool = True,
        progress: bool = True,
        guidance_scale: float = 3.0,
        null_label: int = 0,
        dummy_input_hw: Tuple[int, int] = (224, 224),
        cond=None,
        cond_mask=None,
        key_order: Optional[List[str]] = None,
        class_keys: Optional[Dict[int, List[str]]] = None,
        *,
        noise: Optional[torch.Tensor] = None,
        x_T: Optional[torch.Tensor] = None,
        seed: Optional[int] = None,
        sample_offset: int = 0,
        latent_dump: bool = False,
        return_u8: bool = False,
        sample_steps: Optional[int] = None,
    ):
        """Class + numeric-condition CFG sampling (diff.py:174-369).

        ``sample_steps`` (keyword-only EXTENSION, default None = the reference's full T-step
        ancestral DDPM loop, unchanged): run a deterministic DDIM pass over ``sample_steps``
        evenly strided timesteps instead (must divide T).

        Returns what the reference returns: the latent if ``vae is None``, a (B,3,224,224)
        tensor if ``to_pil=False``, else a list of PIL images.  With ``latent_dump=True`` a
        second value, the (T,B,4,28,28) tensor of x_t before each step, is returned as well.
        ``return_u8=True`` returns the decoded batch as one (B,3,224,224) uint8 device tensor
        (the x*255 / clamp / truncate of reverse_to_img done in the decoder's epilogue).
        """
        device = self.device
        items = _normalise_counts(class_counts)
        y_list: List[int] = []
        for cls, num in items:
            y_list += [cls] * num
        B = len(y_list)
        if any(c < 0 or c > 3 for c in y_list) or not (0 <= int(null_label) <= 3):
            raise IndexError("index out of range in self")     # nn.Embedding(num_classes + 1) in the reference
        y = torch.tensor(y_list, device=device, dtype=torch.long)

        if key_order is None:
            key_order = list(KEY_ORDER)
        if class_keys is None:
            class_keys = {k: list(v) for k, v in CLASS_KEYS.items()}
        vals, msk = _build_condition(cond, cond_mask, y_list, key_order, class_keys, device)

        if z_shape is None:
            if vae is None:
                raise ValueError("z_shape 省略時は vae が必要です。")
            # The reference probes vae.encode(zeros) for the shape (diff.py:315-322); the encoder
            # halves the resolution three times, so the answer is (z_channels, H/8, W/8).
            H, W = dummy_input_hw
            C, Hlat, Wlat = getattr(vae, "z_channels", 4), H // 8, W // 8
        else:
            C, Hlat, Wlat = z_shape
        if K_of(vals) != 12:
            raise ValueError("libdmb200 implements the 12-key condition vector of the reference")

        ctx = self._ctx(model)
        if x_T is None:
            x = torch.randn((B, C, Hlat, Wlat), device=device)
        else:
            x = x_T.to(device=device, dtype=torch.float32).clone()
        if (C, Hlat, Wlat) != (4, 28, 28):
            raise ValueError(f"libdmb200 implements the (4,28,28) latent; got {(C, Hlat, Wlat)}")
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        x, dump = self._run_loop(ctx, x, y, vals, msk, guidance_scale, null_label, progress, noise, seed,
                                 sample_offset, latent_dump, "Sampling (cond+numeric)", sample_steps=sample_steps)

        def ret(v):
            return (v, dump) if latent_dump else v

        if vae is None:
            return ret(x)
        vae.eval()
        if return_u8:
            return ret(vae.decode_u8(x))
        if to_pil and hasattr(vae, "decode_u8"):
            u8 = vae.decode_u8(x).cpu()  # one device->host copy for the whole batch
            return ret([_chw_u8_to_pil(u8[i]) for i in range(B)])
        images = vae.decode(x)
        return ret([self.reverse_to_img(images[i]) for i in range(B)] if to_pil else images)

    # ------------------------------------------------------------------ not on the sampling path
    # ------------------------------------------------------------------ forward process (8f N3)
    def add_noise(self, x_0, t, noise=None):
        """(x_t, noise) like the reference (diff.py:18-30): the same range assertion, the same
        ``torch.randn_like`` draw (pass ``noise`` to supply it), and the blend on the device with
        every operation rounded separately -- bit-identical to the reference for the same noise."""
        from dmb200 import context_for
        assert (t >= 1).all() and (t <= self.num_timesteps).all()
        if noise is None:
            noise = torch.randn_like(x_0)
        ctx = context_for(x_0.device)
        x_t = ctx.add_noise(x_0, t.to(x_0.device), self.alpha_bars.to(x_0.device), noise)
        return x_t, noise

    def denoise(self, model, x, t):
        raise NotImplementedError("unconditional denoise (diff.py:32-56) is outside the CFG sampling path")

    def sample(self, model, x_shape=(20, 3, 80, 80)):
        raise NotImplementedError("pixel-space unconditional sampling (diff.py:66-87) is outside the CFG path")

    def sample_latent(self, model, z_shape=(1000, 4, 28, 28), vae=None, to_pil=True, progress=True):
        raise NotImplementedError("unconditional latent sampling (diff.py:89-125) is outside the CFG path")


# ---------------------------------------------------------------------- helpers
def K_of(vals: torch.Tensor) -> int:
    return int(vals.shape[1])


def _chw_u8_to_pil(x: torch.Tensor):
    from PIL import Image
    arr = x.numpy()
    if arr.shape[0] == 1:
        return Image.fromarray(arr[0], mode="L")
    mode = {3: "RGB", 4: "RGBA"}[arr.shape[0]]
    return Image.fromarray(arr.transpose(1, 2, 0).copy(), mode=mode)


def _normalise_counts(cc) -> List[Tuple[int, int]]:
    """{cls: n} | (cls, n) | [(cls, n), ...] -> [(cls, n)] with n > 0  (diff.py:206-218)."""
    if isinstance(cc, dict):
        pairs = list(cc.items())
    elif isinstance(cc, tuple) and len(cc) == 2:
        pairs = [cc]
    elif isinstance(cc, list):
        pairs = list(cc)
    else:
        raise ValueError("class_counts は {cls:num}, (cls,num), そのリストのいずれか。")
    pairs = [(int(c), int(n)) for c, n in pairs if int(n) > 0]
    if not pairs:
        raise ValueError("生成枚数が0です。")
    return pairs


def _build_condition(cond, cond_mask, y_list, key_order, class_keys, device):
    """(B,K) value / mask tensors from tensor, per-class dict, per-sample list or None inputs,
    with the reference's defaulting rules (diff.py:241-312): a tensor mask defaults to
    ``vals != 0``; dict / list entries set mask 1 unless an explicit mask says otherwise;
    with no condition at all the class's keys are masked in with value 0."""
    B, K = len(y_list), len(key_order)
    pos = {k: i for i, k in enumerate(key_order)}
    vals = cond.to(device) if isinstance(cond, torch.Tensor) else None
    msk = cond_mask.to(device) if isinstance(cond_mask, torch.Tensor) else None

    if vals is not None:
        if vals.ndim != 2 or vals.shape[0] != B or vals.shape[1] != K:
            raise ValueError(f"cond Tensor 形状は (B={B}, K={K}) 必須: got {tuple(vals.shape)}")
        if msk is None:
            msk = (vals != 0).float()
        elif msk.ndim != 2 or msk.shape != vals.shape:
            raise ValueError("cond_mask Tensor 形状は cond と同じ (B,K) 必須。")
        return vals.float(), msk.float()

    v = torch.zeros((B, K), dtype=torch.float32)
    m = torch.zeros((B, K), dtype=torch.float32) if msk is None else msk.detach().float().cpu().clone()

    def put(i, table, explicit):
        for k, val in table.items():
            if k in pos:
                v[i, pos[k]] = float(val)
                m[i, pos[k]] = float(explicit[k]) if (explicit is not None and k in explicit) else 1.0

    if isinstance(cond, dict):
        for i, cls in enumerate(y_list):
            explicit = cond_mask.get(cls) if isinstance(cond_mask, dict) else None
            if cls in cond:
                put(i, cond[cls], explicit)
            if explicit is not None:
                for k, mv in explicit.items():
                    if k in pos:
                        m[i, pos[k]] = float(mv)
    elif isinstance(cond, list):
        if len(cond) != B:
            raise ValueError(f"cond(list) の長さ {len(cond)} が生成枚数 {B} と不一致。")
        for i, table in enumerate(cond):
            explicit = cond_mask[i] if (isinstance(cond_mask, list) and i < len(cond_mask)) else None
            put(i, table, explicit)
        if isinstance(cond_mask, list) and len(cond_mask) == B:
            for i, table in enumerate(cond_mask):
                for k, mv in table.items():
                    if k in pos:
                        m[i, pos[k]] = float(mv)
    else:
        for i, cls in enumerate(y_list):
            for k in class_keys.get(cls, []):
                if k in pos:
                    m[i, pos[k]] = 1.0
    return v.to(device), m.to(device)
