#!/usr/bin/env python
"""Headline benchmark: CFG DDPM sampling throughput (samples/s at T=1000) of the B200-native
path, one process per GPU.

    python bench.py --gpus 1 --steps 3 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference algorithm on the host CPU

A "step" is one pass of the hot path over one batch: T=1000 CFG reverse steps of `--batch`
samples per GPU (two U-Net passes per step as one doubled batch), the VAE decode to u8
images and, for N > 1, the NCCL gather of the images.  Weak scaling: per-GPU batch fixed.
Prints ONE JSON line on rank 0 (see README / DESIGN.md for the keys).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PRODUCT = os.path.join(ROOT, "diffusion-model_b200")
for _p in (ROOT, PRODUCT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "cfg_sampling_throughput_T1000"
UNIT = "samples/s"
UNET_FWD_FLOP = 2_698_887_680     # SURVEY.md 8d, per sample per forward
VAE_DEC_FLOP = 8_819_736_576      # per sample


def flop_per_sample(T):
    """2 x T x U-Net forward + VAE decode; 5,406,595,096,576 at T=1000."""
    return 2 * T * UNET_FWD_FLOP + VAE_DEC_FLOP
# Fallback roofline denominators = the driver's measurement on this pool (MEASURED_PEAKS.json of
# 2026-11-04) in case the file did not travel with the snapshot.
FALLBACK_PEAKS = {"hbm_gbs": 6528.1, "bf16_tflops": 1664.7, "bf16_tflops_sustained": 1362.8}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--batch", type=int, default=256, help="samples per GPU per step")
    ap.add_argument("--T", type=int, default=1000)
    ap.add_argument("--precision", default=os.environ.get("DMB200_PRECISION", "bf16"))
    ap.add_argument("--guidance", type=float, default=3.0)
    ap.add_argument("--cpu-sample-steps", type=int, default=3,
                    help="CFG steps timed on the host CPU for the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            d = json.load(f)
        return {k: float(d[k]) for k in FALLBACK_PEAKS}, "measured (MEASURED_PEAKS.json)"
    except Exception:
        return dict(FALLBACK_PEAKS), "fallback (driver measurement of 2026-11-04, file absent)"


# ---------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [c.strip() for c in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------- CPU baseline
def cpu_reference_sample(batch: int, n_steps: int, T: int, guidance: float):
    """Times the oracle port (oracle/ref_torch.py -- the reference's algorithm on torch CPU fp32;
    the reference itself is Python and cannot travel to the GPU box) on the host cores:
    `n_steps` CFG denoise steps at `batch` plus one VAE decode of 4 (the reference's decode
    chunk, diff.py:353), extrapolated to T steps."""
    import torch
    from oracle import ref_torch as R
    from oracle import synth
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    torch.manual_seed(0)
    sd = UnetCondWithGeomHead().state_dict()
    vsd = VAE().state_dict()
    y, vals, mask = synth.line_conditions(batch, seed=0)
    x_T, noise = synth.noise_bank(n_steps + 1, (batch, 4, 28, 28), seed=1234)
    _, alphas, abars = R.schedule(T)
    x = x_T
    with torch.no_grad():
        t = torch.full((batch,), T, dtype=torch.long)
        x, _ = R.denoise_cond(sd, x, t, y, guidance, 0, vals, mask, noise[0], alphas, abars)  # warm-up
        t0 = time.perf_counter()
        for i in range(n_steps):
            t = torch.full((batch,), T - 1 - i, dtype=torch.long)
            x, _ = R.denoise_cond(sd, x, t, y, guidance, 0, vals, mask, noise[i + 1], alphas, abars)
        step_s = (time.perf_counter() - t0) / n_steps
        nd = min(4, batch)
        R.vae_decode(vsd, x[:nd])
        t0 = time.perf_counter()
        R.vae_decode(vsd, x[:nd])
        dec_s = (time.perf_counter() - t0) / nd
    total_s = T * step_s + batch * dec_s
    return {"value": batch / total_s, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "host_cpus": os.cpu_count(), "step_ms": step_s * 1e3, "decode_ms_per_sample": dec_s * 1e3,
            "sample": f"{n_steps} CFG denoise steps at batch {batch} + VAE decode of {nd}, oracle port "
                      f"(torch CPU fp32), extrapolated to T={T}"}


def run_reference(args):
    """--impl reference: the reference algorithm on the host CPU (oracle port), rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    batch = min(args.batch, 32)  # CPU throughput is flat beyond ~16; keeps the sample bounded
    vals = []
    for _ in range(max(0, args.warmup)):
        cpu_reference_sample(batch, 1, args.T, args.guidance)
    t0 = time.perf_counter()
    for _ in range(max(1, args.steps)):
        vals.append(cpu_reference_sample(batch, args.cpu_sample_steps, args.T, args.guidance))
    wall = time.perf_counter() - t0
    v = statistics.median([r["value"] for r in vals])
    cb = dict(vals[-1]); cb["value"] = v
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"reference algorithm on host CPU: CFG DDPM sampling T={args.T}, batch {batch} line "
                               "conditions, random-init weights; each step = bounded sample extrapolated to T",
                   "batch_per_gpu": batch, "T": args.T, "guidance_scale": args.guidance,
                   "threads": torch.get_num_threads()},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_native(args):
    import torch
    import torch.distributed as dist
    from oracle import synth  # synthetic conditions only (bench input generator)
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    from diff import Diffuser
    from dmb200 import context_for
    from dmb200.sharding import gather_shards, shard_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl native needs a B200: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T = args.batch, args.T
    peaks, peak_src = load_peaks()

    torch.manual_seed(0)
    unet = UnetCondWithGeomHead().eval().to(dev)
    vae = VAE().eval().to(dev)
    unet.precision = vae.precision = args.precision
    ctx = context_for(dev, args.precision)
    ctx.ensure_unet(unet)
    ctx.ensure_vae(vae)
    d = Diffuser(num_timesteps=T, device=dev)
    coef = d.step_coefficients()

    # this rank's shard of the global batch: contiguous, conditions ordered by class globally
    y_all, vals_all, mask_all, _ = synth.mixed_conditions(B * world, seed=0)
    sl = slice(*shard_range(B * world, rank, world))
    y_h, vals_h, mask_h = y_all[sl].contiguous(), vals_all[sl].contiguous(), mask_all[sl].contiguous()
    y, vals, mask = y_h.to(dev), vals_h.to(dev), mask_h.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_pass(step_idx):
        """Device-resident pass: x_T drawn on the device, conditions resident in HBM."""
        x = ctx.philox_normal(B, seed=4321 + step_idx, sample_offset=rank * B, t=0)
        ctx.sample_loop(x, y, 0, vals, mask, args.guidance, coef, T, T, seed=1234 + step_idx,
                        sample_offset=rank * B)
        _, u8 = ctx.vae_decode(x, want_f32=False, want_u8=True)
        if world > 1:
            return gather_shards(u8, world * B)  # one NCCL all-gather of the u8 images
        return u8

    for i in range(args.warmup):
        one_pass(i)
    barrier()
    ctx.reset_launch_count()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        one_pass(100 + i)
    e1.record()
    barrier()
    launches = ctx.launch_count()
    ms_total = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_total = float(ms_total.item())
    value = world * B * args.steps / (ms_total * 1e-3)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    vals_p, mask_p = vals_h.pin_memory(), mask_h.pin_memory()
    xT_p = torch.randn((B, 4, 28, 28)).pin_memory()
    out_p = torch.empty((B, 3, 224, 224), dtype=torch.uint8).pin_memory()
    counts = {}
    for c in y_h.tolist():
        counts[c] = counts.get(c, 0) + 1
    class_counts = [(c, n) for c, n in sorted(counts.items())]

    def e2e_pass():
        v = vals_p.to(dev, non_blocking=True)
        m = mask_p.to(dev, non_blocking=True)
        xt = xT_p.to(dev, non_blocking=True)
        u8 = d.sample_latent_cond(unet, class_counts, vae=vae, progress=False, guidance_scale=args.guidance,
                                  cond=v, cond_mask=m, z_shape=(4, 28, 28), x_T=xt, seed=99,
                                  sample_offset=rank * B, return_u8=True)
        out_p.copy_(u8, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    e2e_steps = max(1, min(args.steps, 2))
    e2e_pass()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(e2e_steps):
        e2e_pass()
    e1.record()
    barrier()
    e2e_ms = torch.tensor([max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)], device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * B * e2e_steps / (float(e2e_ms.item()) * 1e-3)
    clock_info = clocks.stop() if rank == 0 else None
    h2d = (vals_p.numel() + mask_p.numel() + xT_p.numel()) * 4
    d2h = out_p.numel()

    # ---- per-kernel device times of one eager step (CUDA events on the launching stream)
    x = torch.randn((B, 4, 28, 28), device=dev)
    tt = torch.full((B,), T // 2, dtype=torch.long, device=dev)
    ctx.denoise_cond_step(x, tt, y, 0, vals, mask, args.guidance, coef, seed=1)
    torch.cuda.synchronize()
    ctx.set_kernel_timing(True)
    reps = 3
    for _ in range(reps):
        ctx.denoise_cond_step(x, tt, y, 0, vals, mask, args.guidance, coef, seed=1)
    rows = ctx.kernel_timing()
    ctx.set_kernel_timing(False)
    gemm = [r for r in rows if r["label"].startswith(("conv3x3", "linear"))]
    g_ms = sum(r["ms"] for r in gemm)
    g_fl = sum(r["flops"] for r in gemm)
    g_n = sum(r["count"] for r in gemm)
    step_ms_eager = sum(r["ms"] for r in rows) / reps
    achieved = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
    peak = peaks["bf16_tflops_sustained"]
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get("gemm_tc_kernel_dram_bytes_per_launch")
    except Exception:
        pass
    groups = {}
    for r in rows:
        key = r["label"].split(" ")[0]
        g = groups.setdefault(key, {"ms_per_step": 0.0, "launches_per_step": 0})
        g["ms_per_step"] += r["ms"] / reps
        g["launches_per_step"] += r["count"] // reps
    roofline = {
        "kernel": "gemm_tc_kernel (tcgen05 implicit-GEMM: 3x3 convs + attention-block linears)",
        "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
        "peak_source": peak_src + ", sustained bf16 figure (kernel timed inside a long step)",
        "traffic": traffic,
        "flops_per_launch": g_fl / max(g_n, 1), "avg_launch_us": 1e3 * g_ms / max(g_n, 1),
        "launches_per_step": g_n // reps, "share_of_step": g_ms / reps / step_ms_eager,
        "whole_job": {"achieved": value / world * flop_per_sample(T) / 1e12, "unit": "TFLOP/s",
                      "frac": value / world * flop_per_sample(T) / 1e12 / peak,
                      "note": "samples/s/GPU x (2 T x 2.6989 GFLOP + 8.82 GFLOP) per sample (SURVEY.md 8d)"},
        "kernel_groups_ms_per_denoise_step": {k: round(v["ms_per_step"], 4) for k, v in groups.items()},
    }

    # ---- BASELINE's second figure: batch-1 ms per denoise step (config 5, the generate_steps
    # latency path): the CUDA-graph loop on one sample, with and without the per-step latent dump
    batch1 = None
    if rank == 0:
        y1, v1, m1 = y[:1].contiguous(), vals[:1].contiguous(), mask[:1].contiguous()
        n1 = min(T, 200)
        res = {}
        for dump in (False, True):
            xd = torch.empty((n1, 1, 4, 28, 28), device=dev) if dump else None
            x1 = ctx.philox_normal(1, seed=7, sample_offset=0, t=0)
            ctx.sample_loop(x1, y1, 0, v1, m1, args.guidance, coef, T, n1, seed=5, x_dump=xd)
            torch.cuda.synchronize()
            x1 = ctx.philox_normal(1, seed=8, sample_offset=0, t=0)
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b0.record()
            ctx.sample_loop(x1, y1, 0, v1, m1, args.guidance, coef, T, n1, seed=6, x_dump=xd)
            b1.record()
            torch.cuda.synchronize()
            res["with_latent_dump" if dump else "plain"] = b0.elapsed_time(b1) / n1
        batch1 = {"ms_per_denoise_step": res["plain"], "ms_per_denoise_step_with_latent_dump": res["with_latent_dump"],
                  "steps_timed": n1, "note": "batch 1, CFG pair = 2 U-Net passes per step, CUDA-graph loop"}

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_reference_sample(min(B, 16), args.cpu_sample_steps, T, args.guidance)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "tf32",
            "data": "synthetic",
            "config": {
                "workload": f"CFG DDPM sampling T={T}, batch {B} per GPU of mixed line/circle/arc conditions "
                            f"(cond_vals+cond_mask), guidance {args.guidance}, + VAE decode to u8 "
                            "(BASELINE config 3; config 4 sharding for n_gpus > 1); random-init weights seed 0",
                "batch_per_gpu": B, "global_batch": B * world, "T": T, "guidance_scale": args.guidance,
                "parallelism": f"sample-sharded x{world}, NCCL all-gather of u8 images" if world > 1 else "single GPU",
                "l2_policy": "working set per denoise step (activations of the doubled batch, several GB at "
                             "batch 256) exceeds the 126 MB L2; no explicit flush",
                "noise": "device Philox keyed by (seed, global sample index, t)",
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps,
                    "api": "Diffuser.sample_latent_cond(..., return_u8=True) with pinned host cond/x_T in, "
                           "pinned host u8 images out"},
            "gpu_launches": int(launches),
            "clocks": clock_info,
            "roofline": roofline,
            "cpu_baseline": cpu,
            "ms_per_denoise_step": ms_total / args.steps / T,
            "batch1": batch1,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
