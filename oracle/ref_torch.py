"""CPU fp32 restatement of the reference's conditional reverse-diffusion path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Every function is a
functional restatement over a flat ``state_dict`` in the reference's own key
layout; nothing here builds ``nn.Module`` objects, and nothing here is imported
by the product.  The arithmetic lives in torch's ATen CPU kernels
(torch 2.11.0+cu128 in this image; the reference pins no version,
requirements.yml:6-7).

Citations are ``file:line`` relative to ``/root/reference``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]

KEY_ORDER = ["x1", "y1", "x2", "y2", "cx", "cy", "cr", "ax", "ay", "ar", "theta1", "theta2"]
CLASS_KEYS = {1: KEY_ORDER[0:4], 2: KEY_ORDER[4:7], 3: KEY_ORDER[7:12]}
VAE_SCALE = 0.18215


# --------------------------------------------------------------------------
# U-Net building blocks
# --------------------------------------------------------------------------
def res_block(sd: SD, p: str, x: torch.Tensor, residual: bool) -> torch.Tensor:
    """ResBlock: conv3x3 -> GroupNorm(1) -> GELU(erf) -> conv3x3 -> GroupNorm(1),
    optionally GELU(x + .).  models/unet_cond.py:10-30."""
    h = F.conv2d(x, sd[f"{p}.double_conv.0.weight"], None, padding=1)
    h = F.group_norm(h, 1, sd[f"{p}.double_conv.1.weight"], sd[f"{p}.double_conv.1.bias"], 1e-5)
    h = F.gelu(h)
    h = F.conv2d(h, sd[f"{p}.double_conv.3.weight"], None, padding=1)
    h = F.group_norm(h, 1, sd[f"{p}.double_conv.4.weight"], sd[f"{p}.double_conv.4.bias"], 1e-5)
    return F.gelu(x + h) if residual else h


def attention_block(sd: SD, p: str, x: torch.Tensor, heads: int = 4) -> torch.Tensor:
    """AttenionBlock (sic): tokens = pixels; the residual base is the layer-normed
    tensor.  models/unet_cond.py:32-52.  nn.MultiheadAttention (batch_first,
    4 heads, in/out-proj bias, scale 1/sqrt(d_h)) is written out explicitly."""
    B, C, H, W = x.shape
    L = H * W
    d = C // heads
    tok = x.reshape(B, C, L).transpose(1, 2)
    xl = F.layer_norm(tok, (C,), sd[f"{p}.ln.weight"], sd[f"{p}.ln.bias"], 1e-5)
    qkv = F.linear(xl, sd[f"{p}.mha.in_proj_weight"], sd[f"{p}.mha.in_proj_bias"])
    q, k, v = qkv.chunk(3, dim=-1)

    def split(z):
        return z.reshape(B, L, heads, d).transpose(1, 2)

    q, k, v = split(q), split(k), split(v)
    s = torch.matmul(q, k.transpose(-1, -2)) * (1.0 / math.sqrt(d))
    a = torch.matmul(torch.softmax(s, dim=-1), v)
    a = a.transpose(1, 2).reshape(B, L, C)
    a = F.linear(a, sd[f"{p}.mha.out_proj.weight"], sd[f"{p}.mha.out_proj.bias"]) + xl
    h = F.layer_norm(a, (C,), sd[f"{p}.ff_self.0.weight"], sd[f"{p}.ff_self.0.bias"], 1e-5)
    h = F.linear(h, sd[f"{p}.ff_self.1.weight"], sd[f"{p}.ff_self.1.bias"])
    h = F.linear(F.gelu(h), sd[f"{p}.ff_self.3.weight"], sd[f"{p}.ff_self.3.bias"])
    out = h + a
    return out.transpose(1, 2).reshape(B, C, H, W)


def _emb_proj(sd: SD, p: str, emb: torch.Tensor) -> torch.Tensor:
    """emb_layer = SiLU -> Linear(256, C_out).  models/unet_cond.py:62-65, 82-85."""
    return F.linear(F.silu(emb), sd[f"{p}.emb_layer.1.weight"], sd[f"{p}.emb_layer.1.bias"])


def down(sd: SD, p: str, x: torch.Tensor, emb: torch.Tensor) -> torch.Tensor:
    """Down: MaxPool2d(2) (floor) -> RB(residual) -> RB -> + emb.  unet_cond.py:54-70."""
    x = F.max_pool2d(x, 2)
    x = res_block(sd, f"{p}.maxpool_conv.1", x, True)
    x = res_block(sd, f"{p}.maxpool_conv.2", x, False)
    return x + _emb_proj(sd, p, emb)[:, :, None, None]


def up(sd: SD, p: str, x: torch.Tensor, skip: torch.Tensor, emb: torch.Tensor) -> torch.Tensor:
    """Up: bilinear x2 (align_corners) -> zero-pad right/bottom to the skip size ->
    cat([skip, x]) -> RB(residual) -> RB(mid = in/2) -> + emb.  unet_cond.py:72-100."""
    x = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)
    dy = skip.size(2) - x.size(2)
    dx = skip.size(3) - x.size(3)
    if dy != 0 or dx != 0:
        x = F.pad(x, [max(0, dx // 2), max(0, dx - dx // 2), max(0, dy // 2), max(0, dy - dy // 2)])
    x = torch.cat([skip, x], dim=1)
    x = res_block(sd, f"{p}.conv.0", x, True)
    x = res_block(sd, f"{p}.conv.1", x, False)
    return x + _emb_proj(sd, p, emb)[:, :, None, None]


def pos_encoding(t: torch.Tensor, channels: int = 256) -> torch.Tensor:
    """[sin block | cos block] of t * 10000^(-2j/channels).  unet_cond.py:155-161.
    ``t`` is int64 (B,); the product t*inv_freq is formed in fp32."""
    inv_freq = 1.0 / (10000 ** (torch.arange(0, channels, 2).float() / channels))
    arg = t.unsqueeze(-1).repeat(1, channels // 2) * inv_freq
    return torch.cat([torch.sin(arg), torch.cos(arg)], dim=-1)


def embedding(sd: SD, t: torch.Tensor, y: torch.Tensor,
              cond_vals: Optional[torch.Tensor], cond_mask: Optional[torch.Tensor]) -> torch.Tensor:
    """fused_embedding + cond_mlp.  unet_cond.py:163-167; unet_cond_geom.py:89-95."""
    emb = pos_encoding(t) + F.embedding(y, sd["class_emb.weight"])
    if cond_vals is not None and cond_mask is not None:
        feat = torch.cat([cond_vals, cond_mask], dim=1)
        h = F.silu(F.linear(feat, sd["cond_mlp.0.weight"], sd["cond_mlp.0.bias"]))
        emb = emb + F.linear(h, sd["cond_mlp.2.weight"], sd["cond_mlp.2.bias"])
    return emb


def unet_trunk(sd: SD, x: torch.Tensor, emb: torch.Tensor,
               taps: Optional[Dict[str, torch.Tensor]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """unet_forward_with_feat.  unet_cond_geom.py:52-76.  ``taps`` (optional dict)
    receives named intermediate activations for layer-level parity tests."""
    def rec(name, v):
        if taps is not None:
            taps[name] = v
        return v

    x1 = rec("inc", res_block(sd, "inc", x, False))
    x2 = rec("down1", down(sd, "down1", x1, emb))
    x2 = rec("sa1", attention_block(sd, "sa1", x2))
    x3 = rec("down2", down(sd, "down2", x2, emb))
    x3 = rec("sa2", attention_block(sd, "sa2", x3))
    x4 = rec("down3", down(sd, "down3", x3, emb))
    x4 = rec("sa3", attention_block(sd, "sa3", x4))
    x4 = rec("bot1", res_block(sd, "bot1", x4, False))
    x4 = rec("bot2", res_block(sd, "bot2", x4, False))
    x4 = rec("bot3", res_block(sd, "bot3", x4, False))
    h = rec("up1", up(sd, "up1", x4, x3, emb))
    h = rec("sa4", attention_block(sd, "sa4", h))
    h = rec("up2", up(sd, "up2", h, x2, emb))
    h = rec("sa5", attention_block(sd, "sa5", h))
    h = rec("up3", up(sd, "up3", h, x1, emb))
    h = rec("sa6", attention_block(sd, "sa6", h))
    eps = F.conv2d(h, sd["out.weight"], sd["out.bias"])
    return eps, h


def unet_forward(sd: SD, x: torch.Tensor, t: torch.Tensor, y: torch.Tensor,
                 cond_vals: Optional[torch.Tensor] = None, cond_mask: Optional[torch.Tensor] = None,
                 taps: Optional[Dict[str, torch.Tensor]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """UnetCondWithGeomHead.forward -> (eps, geom).  unet_cond_geom.py:79-100, 8-23."""
    emb = embedding(sd, t, y, cond_vals, cond_mask)
    if taps is not None:
        taps["emb"] = emb
    eps, feat = unet_trunk(sd, x, emb, taps)
    g = feat.mean(dim=(2, 3))
    g = F.silu(F.linear(g, sd["geom_head.mlp.0.weight"], sd["geom_head.mlp.0.bias"]))
    geom = F.linear(g, sd["geom_head.mlp.2.weight"], sd["geom_head.mlp.2.bias"])
    return eps, geom


# --------------------------------------------------------------------------
# Diffusion process
# --------------------------------------------------------------------------
def schedule(num_timesteps: int = 1000, beta_start: float = 1e-4, beta_end: float = 0.02):
    """Linear-beta DDPM schedule in fp32.  diff.py:11-16."""
    betas = torch.linspace(beta_start, beta_end, num_timesteps)
    alphas = 1 - betas
    alpha_bars = torch.cumprod(alphas, dim=0)
    return betas, alphas, alpha_bars


def ddpm_update(x: torch.Tensor, eps: torch.Tensor, t: torch.Tensor, noise: torch.Tensor,
                alphas: torch.Tensor, alpha_bars: torch.Tensor) -> torch.Tensor:
    """Posterior mean + sigma*noise with noise zeroed at t == 1.  diff.py:141-144, 158-162."""
    t_idx = t - 1
    alpha = alphas[t_idx].view(-1, 1, 1, 1)
    alpha_bar = alpha_bars[t_idx].view(-1, 1, 1, 1)
    alpha_bar_prev = alpha_bars[torch.clamp(t_idx - 1, min=0)].view(-1, 1, 1, 1)
    noise = noise.clone()
    noise[t == 1] = 0
    mu = (x - ((1 - alpha) / torch.sqrt(1 - alpha_bar)) * eps) / torch.sqrt(alpha)
    std = torch.sqrt((1 - alpha) * (1 - alpha_bar_prev) / (1 - alpha_bar))
    return mu + noise * std


def cfg_eps(sd: SD, x, t, y, guidance_scale: float, null_label: int, cond_vals, cond_mask):
    """eps_u + g (eps_c - eps_u); the null pass keeps cond_vals/cond_mask.  diff.py:147-151."""
    y_null = torch.full_like(y, null_label)
    eps_u, _ = unet_forward(sd, x, t, y_null, cond_vals, cond_mask)
    eps_c, _ = unet_forward(sd, x, t, y, cond_vals, cond_mask)
    return eps_u + guidance_scale * (eps_c - eps_u), eps_u, eps_c


def denoise_cond(sd: SD, x, t, y, guidance_scale, null_label, cond_vals, cond_mask,
                 noise, alphas, alpha_bars):
    """One CFG reverse step with injected noise.  diff.py:127-162."""
    T = alphas.numel()
    assert bool((t >= 1).all()) and bool((t <= T).all())
    eps, _, _ = cfg_eps(sd, x, t, y, guidance_scale, null_label, cond_vals, cond_mask)
    return ddpm_update(x, eps, t, noise, alphas, alpha_bars), eps


def sample_latent(sd: SD, x_T: torch.Tensor, y: torch.Tensor, cond_vals, cond_mask,
                  noise_fn, num_timesteps: int = 1000, guidance_scale: float = 3.0,
                  null_label: int = 0, record: Optional[list] = None) -> torch.Tensor:
    """The T-step loop of sample_latent_cond.  diff.py:327-344.
    ``noise_fn(i)`` returns the (B,4,28,28) noise for step ``i`` (i = T..1)."""
    _, alphas, alpha_bars = schedule(num_timesteps)
    x = x_T
    B = x.shape[0]
    for i in range(num_timesteps, 0, -1):
        t = torch.full((B,), i, dtype=torch.long)
        x, eps = denoise_cond(sd, x, t, y, guidance_scale, null_label, cond_vals, cond_mask,
                              noise_fn(i), alphas, alpha_bars)
        if record is not None:
            record.append((i, eps, x))
    return x


def ddim_update(x: torch.Tensor, eps: torch.Tensor, t: int, t_prev: int, alpha_bars: torch.Tensor) -> torch.Tensor:
    """EXTENSION, not in the reference: deterministic DDIM step (Song et al.
    2021, eq. 12 with sigma = 0) from timestep t to t_prev (0 = the clean sample), in float64:
    x0 = (x - sqrt(1-ab_t) eps) / sqrt(ab_t);  x' = sqrt(ab_prev) x0 + sqrt(1-ab_prev) eps."""
    ab = alpha_bars.double()
    ab_t = ab[t - 1]
    ab_p = ab[t_prev - 1] if t_prev >= 1 else torch.tensor(1.0, dtype=torch.float64)
    x0 = (x.double() - torch.sqrt(1 - ab_t) * eps.double()) / torch.sqrt(ab_t)
    return (torch.sqrt(ab_p) * x0 + torch.sqrt(1 - ab_p) * eps.double()).float()


def sample_latent_strided(sd: SD, x_T: torch.Tensor, y: torch.Tensor, cond_vals, cond_mask, sample_steps: int,
                          num_timesteps: int = 1000, guidance_scale: float = 3.0, null_label: int = 0,
                          record: Optional[list] = None) -> torch.Tensor:
    """EXTENSION: the CFG loop over T, T-s, ..., s (s = T / sample_steps) with ddim_update."""
    _, _, alpha_bars = schedule(num_timesteps)
    assert num_timesteps % sample_steps == 0
    s = num_timesteps // sample_steps
    x = x_T
    B = x.shape[0]
    for i in range(num_timesteps, 0, -s):
        t = torch.full((B,), i, dtype=torch.long)
        eps, _, _ = cfg_eps(sd, x, t, y, guidance_scale, null_label, cond_vals, cond_mask)
        x = ddim_update(x, eps, i, i - s, alpha_bars)
        if record is not None:
            record.append((i, eps, x))
    return x


# --------------------------------------------------------------------------
# VAE decoder and image conversion
# --------------------------------------------------------------------------
def vae_decode(vsd: SD, z: torch.Tensor) -> torch.Tensor:
    """VAE.decode: z/0.18215 -> dec -> sigmoid.  models/vae.py:35-49, 64-69."""
    h = z / VAE_SCALE
    gn = lambda h, i: F.gelu(F.group_norm(h, 8, vsd[f"dec.{i}.weight"], vsd[f"dec.{i}.bias"], 1e-5))
    h = gn(F.conv2d(h, vsd["dec.0.weight"], vsd["dec.0.bias"], padding=1), 1)
    h = gn(F.conv_transpose2d(h, vsd["dec.3.weight"], vsd["dec.3.bias"], stride=2, padding=1), 4)
    h = gn(F.conv2d(h, vsd["dec.6.weight"], vsd["dec.6.bias"], padding=1), 7)
    h = gn(F.conv_transpose2d(h, vsd["dec.9.weight"], vsd["dec.9.bias"], stride=2, padding=1), 10)
    h = gn(F.conv2d(h, vsd["dec.12.weight"], vsd["dec.12.bias"], padding=1), 13)
    h = gn(F.conv_transpose2d(h, vsd["dec.15.weight"], vsd["dec.15.bias"], stride=2, padding=1), 16)
    h = F.conv2d(h, vsd["dec.18.weight"], vsd["dec.18.bias"], padding=1)
    return torch.sigmoid(h)


def vae_encode(vsd: SD, x: torch.Tensor, eps: torch.Tensor, taps: Optional[Dict[str, torch.Tensor]] = None):
    """VAE.encode with the normal draw passed in: (z, kl_mean, mu, logvar_clamped).
    models/vae.py:16-31 (enc), 51-62 (heads, clamp, reparameterisation, KL).  taps (optional)
    receives the outputs of enc[i] under "venc.<i>" (i = index in the nn.Sequential)."""
    def rec(i, v):
        if taps is not None:
            taps[f"venc.{i}"] = v
        return v
    gn = lambda h, i: rec(i + 1, F.gelu(F.group_norm(h, 8, vsd[f"enc.{i}.weight"], vsd[f"enc.{i}.bias"], 1e-5)))
    conv = lambda h, i, stride: rec(i, F.conv2d(h, vsd[f"enc.{i}.weight"], vsd[f"enc.{i}.bias"], stride=stride, padding=1))
    h = gn(conv(x, 0, 1), 1)
    h = gn(conv(h, 3, 2), 4)
    h = gn(conv(h, 6, 1), 7)
    h = gn(conv(h, 9, 2), 10)
    h = gn(conv(h, 12, 1), 13)
    h = gn(conv(h, 15, 2), 16)
    mu = F.conv2d(h, vsd["to_mu.weight"], vsd["to_mu.bias"])
    logvar = F.conv2d(h, vsd["to_logvar.weight"], vsd["to_logvar.bias"]).clamp(-30.0, 20.0)
    std = torch.exp(0.5 * logvar)
    z = (mu + eps * std) * VAE_SCALE
    kl = 0.5 * torch.sum(torch.exp(logvar) + mu ** 2 - 1.0 - logvar, dim=(1, 2, 3)) / (x.size(2) * x.size(3))
    return z, kl.mean(), mu, logvar


def add_noise(x_0: torch.Tensor, t: torch.Tensor, noise: torch.Tensor, alpha_bars: torch.Tensor) -> torch.Tensor:
    """Diffuser.add_noise with the normal draw passed in.  diff.py:18-30."""
    ab = alpha_bars[t - 1].view(-1, 1, 1, 1)
    return torch.sqrt(ab) * x_0 + torch.sqrt(1 - ab) * noise


def to_u8(img: torch.Tensor) -> torch.Tensor:
    """x*255 -> clamp(0,255) -> uint8 (truncation).  diff.py:58-61."""
    return (img * 255).clamp(0, 255).to(torch.uint8)


def u8_to_gray(u8_chw: np.ndarray) -> np.ndarray:
    """PIL ``convert("L")`` of an RGB u8 image (ITU-R 601-2 luma, integer form used by
    Pillow: (19595 R + 38470 G + 7471 B + 0x8000) >> 16).  eval_iou_noise.py:87."""
    r, g, b = (u8_chw[i].astype(np.uint32) for i in range(3))
    return ((19595 * r + 38470 * g + 7471 * b + 0x8000) >> 16).astype(np.uint8)


def binarise(u8_chw: np.ndarray, threshold: int = 128) -> np.ndarray:
    """load_binary_mask with invert=True: dark pixels are foreground.  eval_iou_noise.py:77-94."""
    return u8_to_gray(u8_chw) < threshold


# --------------------------------------------------------------------------
# Scoring (eval_iou_noise.compute_metrics)
# --------------------------------------------------------------------------
def compute_metrics(gt: np.ndarray, pred: np.ndarray, sigma: float = 2.0) -> Dict[str, float]:
    """IoU, GT-IoU, far-noise ratio and Gaussian-weighted recall of two boolean masks.
    eval_iou_noise.py:162-272 (scipy EDT backend)."""
    from scipy.ndimage import distance_transform_edt
    if gt.shape != pred.shape:
        raise ValueError(f"Shape mismatch: gt{gt.shape} vs pred{pred.shape}")
    inter = int(np.logical_and(gt, pred).sum(dtype=np.int64))
    union = int(np.logical_or(gt, pred).sum(dtype=np.int64))
    gt_area = int(gt.sum(dtype=np.int64))
    pred_area = int(pred.sum(dtype=np.int64))
    iou = inter / union if union > 0 else 1.0
    gt_iou = inter / gt_area if gt_area > 0 else 1.0
    dist = distance_transform_edt(~gt).astype(np.float64) if (gt_area or pred_area) else None
    if pred_area == 0:
        far = 0.0
    else:
        far = float(np.logical_and(pred, dist > sigma).sum(dtype=np.int64) / pred_area)
    if gt_area == 0:
        gauss = 1.0
    else:
        w = np.exp(-(dist ** 2) / (2.0 * sigma ** 2))
        gauss = float((pred.astype(np.float64) * w).sum(dtype=np.float64) / gt_area)
    return {"iou": float(iou), "gt_iou": float(gt_iou), "far_noise_ratio": far,
            "gauss_recall": gauss, "inter": float(inter), "union": float(union),
            "gt_area": float(gt_area), "pred_area": float(pred_area),
            "fp": float(np.logical_and(pred, ~gt).sum(dtype=np.int64))}


# --------------------------------------------------------------------------
# Condition front-end (EntityCsvSampler)
# --------------------------------------------------------------------------
def norm_angle(v: np.ndarray) -> np.ndarray:
    """|v| > 1 is degrees -> (v mod 360)/360, else unchanged.  entityCsvSampler.py:192-199."""
    out = v.astype(np.float32).copy()
    deg = np.abs(out) > 1.0
    out[deg] = (out[deg] % 360.0) / 360.0
    return out


def build_vals_mask(rows: np.ndarray, class_id: int,
                    base_wh: Tuple[float, float] = (400, 400)) -> Tuple[np.ndarray, np.ndarray]:
    """13-column CSV rows -> (vals, mask) (N,12) float32.  entityCsvSampler.py:101-163.
    x/drawW, 1 - y/drawH (Y flip), r/drawW; mask = 1 on the class's keys."""
    rows = np.asarray(rows)
    W, H = np.float32(base_wh[0]), np.float32(base_wh[1])
    n = rows.shape[0]
    vals = np.zeros((n, 12), np.float32)
    mask = np.zeros((n, 12), np.float32)
    col = lambda j: rows[:, j].astype(np.float32)
    nx = lambda a: a / W
    ny = lambda a: (1.0 - (a / H)).astype(np.float32)
    if class_id == 1:
        vals[:, 0], vals[:, 1], vals[:, 2], vals[:, 3] = nx(col(1)), ny(col(2)), nx(col(3)), ny(col(4))
        mask[:, 0:4] = 1.0
    elif class_id == 2:
        vals[:, 4], vals[:, 5], vals[:, 6] = nx(col(5)), ny(col(6)), nx(col(7))
        mask[:, 4:7] = 1.0
    elif class_id == 3:
        vals[:, 7], vals[:, 8], vals[:, 9] = nx(col(8)), ny(col(9)), nx(col(10))
        vals[:, 10], vals[:, 11] = norm_angle(col(11)), norm_angle(col(12))
        mask[:, 7:12] = 1.0
    else:
        raise ValueError("class_id must be 1(line), 2(circle), or 3(arc).")
    return vals, mask
