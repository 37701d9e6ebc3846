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
    ap.add_argument("--config4-batch", type=int, default=4096,
                    help="global batch of the BASELINE config-4 block (sharded over the ranks); 0 skips it")
    ap.add_argument("--config4-chunk", type=int, default=512, help="samples per plan launch inside config 4")
    ap.add_argument("--cpu-leg", default=None, help=argparse.SUPPRESS)   # internal: run one CPU leg and print its JSON
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the tf32 block, the batch-1 eager figure and the config-1 CPU run")
    return ap.parse_args()


def host_threads() -> int:
    """All the host threads this process may use.  torchrun exports OMP_NUM_THREADS=1 to its
    workers; the CPU legs set the torch thread count explicitly so they do not collapse to one."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


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
def cpu_env():
    """Environment for a CPU leg: every host thread for OpenMP / MKL whatever the launcher exported
    (torchrun sets OMP_NUM_THREADS=1), and no GPU."""
    env = dict(os.environ)
    n = str(host_threads())
    env.update(OMP_NUM_THREADS=n, MKL_NUM_THREADS=n, CUDA_VISIBLE_DEVICES="", DMB_BENCH_CPU_ENV="1")
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID",
              "GROUP_RANK", "LOCAL_WORLD_SIZE", "ROLE_RANK", "ROLE_WORLD_SIZE"):
        env.pop(k, None)
    return env


def cpu_leg(kind: str, **kw):
    """Runs one CPU leg (`baseline` = cpu_reference_sample, `config1` = cpu_config1) in a FRESH process:
    inside a process that holds a CUDA context and an NCCL communicator under torchrun the OpenMP
    pool, created with one thread, ran the same code 100x slower even after set_num_threads
    (measured: 1794 vs 17 ms per batch-1 step on a 24-thread host)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--cpu-leg", json.dumps({"kind": kind, **kw})]
    out = subprocess.run(cmd, env=cpu_env(), capture_output=True, text=True, timeout=1200)
    if out.returncode != 0:
        raise RuntimeError(f"CPU leg {kind} failed: {out.stderr[-2000:]}")
    return json.loads([ln for ln in out.stdout.splitlines() if ln.startswith("{")][-1])


def run_cpu_leg(spec: str):
    kw = json.loads(spec)
    kind = kw.pop("kind")
    fn = {"baseline": cpu_reference_sample, "config1": cpu_config1}[kind]
    print(json.dumps(fn(**kw)), flush=True)


def cpu_reference_sample(batch: int, n_steps: int, T: int, guidance: float):
    """Times the oracle port (oracle/ref_torch.py -- the reference's algorithm on torch CPU fp32;
    the reference itself is Python and cannot travel to the GPU box) on ALL the host threads:
    `n_steps` CFG denoise steps at `batch` mixed line/circle/arc conditions (the native arm's
    recipe) plus one VAE decode of 4 (the reference's decode chunk, diff.py:353), extrapolated to
    T steps."""
    import torch
    from oracle import ref_torch as R
    from oracle import synth
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    torch.set_num_threads(host_threads())
    torch.manual_seed(0)
    sd = UnetCondWithGeomHead().state_dict()
    vsd = VAE().state_dict()
    y, vals, mask, _ = synth.mixed_conditions(batch, seed=0)
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
            "host_cpus": os.cpu_count(), "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
            "step_ms": step_s * 1e3, "decode_ms_per_sample": dec_s * 1e3,
            "sample": f"{n_steps} CFG denoise steps at batch {batch} (mixed line/circle/arc) + VAE decode of {nd}, "
                      f"oracle port (torch CPU fp32, {torch.get_num_threads()} threads), extrapolated to T={T}"}


def cpu_config1(T: int, guidance: float, n_steps=None):
    """BASELINE config 1: batch 1, one line condition, the reference's algorithm on the host CPU.
    n_steps None = the whole T-step trajectory run outright (about a minute),
    otherwise `n_steps` consecutive steps after two warm-up steps."""
    import torch
    from oracle import ref_torch as R
    from oracle import synth
    from models.unet_cond_geom import UnetCondWithGeomHead
    from models.vae import VAE
    torch.set_num_threads(host_threads())
    torch.manual_seed(0)
    sd = UnetCondWithGeomHead().state_dict()
    vsd = VAE().state_dict()
    y, vals, mask = synth.line_conditions(1, seed=0)
    n = T if n_steps is None else n_steps
    x_T, noise = synth.noise_bank(n + 2, (1, 4, 28, 28), seed=1234)
    _, alphas, abars = R.schedule(T)
    x = x_T
    with torch.no_grad():
        for i in range(2):
            R.denoise_cond(sd, x, torch.full((1,), T, dtype=torch.long), y, guidance, 0, vals, mask, noise[i], alphas, abars)
        t0 = time.perf_counter()
        for i in range(n):
            t = torch.full((1,), T - i, dtype=torch.long)
            x, _ = R.denoise_cond(sd, x, t, y, guidance, 0, vals, mask, noise[i + 2], alphas, abars)
        loop_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        R.vae_decode(vsd, x)
        dec_s = time.perf_counter() - t0
    total = loop_s * (T / n) + dec_s
    return {"ms_per_denoise_step": 1e3 * loop_s / n, "samples_per_s": 1.0 / total, "decode_ms": dec_s * 1e3,
            "steps_run": n, "T": T, "cores": torch.get_num_threads(), "kind": "port",
            "sample": ("whole trajectory run outright" if n_steps is None else f"{n} steps, extrapolated to T={T}")
                      + f", batch 1, one line condition, oracle port on {torch.get_num_threads()} host threads"}


def run_reference(args):
    """--impl reference: the reference algorithm on the host CPU (oracle port), rank 0 only, on all
    host threads whatever the launcher exported."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if os.environ.get("DMB_BENCH_CPU_ENV") != "1":
        # same process image, clean environment (all host threads, no GPU, no rendezvous variables)
        sys.stdout.flush()
        os.execve(sys.executable, [sys.executable, os.path.abspath(__file__)] + sys.argv[1:], cpu_env())
    import torch
    torch.set_num_threads(host_threads())
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
    config1 = None
    if not args.no_extras:
        # config 1 outright at N=1 (about a minute); a bounded sample when launched for N > 1
        config1 = cpu_config1(args.T, args.guidance, None if args.gpus == 1 else 20)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"reference algorithm on host CPU: CFG DDPM sampling T={args.T}, batch {batch} = "
                               f"min(native batch {args.batch}, 32) of mixed line/circle/arc conditions "
                               "(the native arm's recipe), random-init weights seed 0; each step = "
                               f"{args.cpu_sample_steps} CFG steps + decode, extrapolated to T",
                   "batch_per_gpu": batch, "T": args.T, "guidance_scale": args.guidance,
                   "threads": torch.get_num_threads(), "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS")},
        "cpu_baseline": cb,
        "config1_batch1": config1,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------- BASELINE config 4
def run_config4(args, ctx, dev, world, rank, T, coef, barrier):
    """BASELINE config 4: a global batch of 4096 synthetic conditions sharded contiguously over the
    ranks (4096 / N per GPU, run as plan launches of `--config4-chunk` samples), the u8 images
    gathered with ONE NCCL all-gather, then (rank 0, outside the timed region):
      gather_check     whole chunks that OTHER ranks produced are recomputed on rank 0 and compared
                       byte for byte with the gathered images (noise is keyed by the global sample
                       index, so the N-rank result must equal the 1-rank result);
      eval_iou_parity  eval_iou_noise scoring of all gathered images on the device
                       (dmb_score_masks) against a deterministic pseudo ground truth rendered from
                       the conditions, with the oracle's compute_metrics on 16 of them as the checker.
    Returns the block for the JSON line (rank 0) or None."""
    import random
    import numpy as np
    import torch
    import torch.distributed as dist
    from oracle import synth
    from dmb200.sharding import gather_shards, shard_range
    G, chunk = args.config4_batch, args.config4_chunk
    y_all, vals_all, mask_all, counts = synth.mixed_conditions(G, seed=4)
    yd, vd, md = y_all.to(dev), vals_all.to(dev), mask_all.to(dev)
    b, e = shard_range(G, rank, world)

    def run_chunk(g0, n, steps=T, t_start=T):
        x = ctx.philox_normal(n, seed=777, sample_offset=g0, t=0)
        ctx.sample_loop(x, yd[g0:g0 + n], 0, vd[g0:g0 + n], md[g0:g0 + n], args.guidance, coef, t_start, steps,
                        seed=888, sample_offset=g0)
        _, u8 = ctx.vae_decode(x, want_f32=False, want_u8=True)
        return u8

    run_chunk(b, min(chunk, e - b), steps=min(20, T))  # plan + graph + decoder warm-up
    import torch.cuda as tc
    e0, e1 = tc.Event(enable_timing=True), tc.Event(enable_timing=True)
    barrier()
    e0.record()
    parts = [run_chunk(g0, min(chunk, e - g0)) for g0 in range(b, e, chunk)]
    local = parts[0] if len(parts) == 1 else torch.cat(parts, 0)
    full = gather_shards(local, G) if world > 1 else local
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    if rank != 0:
        return None
    block = {"global_batch": G, "batch_per_gpu": e - b, "chunk": chunk, "n_gpus": world, "T": T, "ms": ms,
             "samples_per_s": G / (ms * 1e-3), "scaling": "strong (total work fixed at 4096 samples)",
             "collective": "one all_gather_into_tensor of (4096/N, 3, 224, 224) u8 per rank" if world > 1 else "none"}
    # ---- gather_check: recompute chunks of other ranks (any chunk at N=1: determinism)
    starts = [g0 for g0 in range(0, G, chunk) if world == 1 or not (b <= g0 < e)]
    pick = sorted(random.Random(0).sample(starts, min(2, len(starts))))
    equal, n_checked = True, 0
    for g0 in pick:
        n = min(chunk, G - g0)
        again = run_chunk(g0, n)
        equal = equal and bool(torch.equal(again, full[g0:g0 + n]))
        n_checked += n
    block["gather_check"] = {"n": n_checked, "global_chunk_starts": pick, "equal": equal,
                             "what": "chunks produced by other ranks recomputed on rank 0, byte equality with "
                                     "the gathered u8 images" if world > 1 else
                                     "chunks recomputed on the same GPU (determinism), byte equality"}
    # ---- eval_iou_parity
    import eval_iou_noise as E
    from oracle import ref_torch as R
    gt = synth.render_gt(y_all.numpy(), vals_all.numpy())
    gt_dev = torch.from_numpy(gt).to(dev)[:, None].contiguous()
    E.compute_metrics_batch(gt_dev[:8], full[:8])
    tc.synchronize()
    t0 = time.perf_counter()
    metrics = E.compute_metrics_batch(gt_dev, full, sigma=2.0)
    tc.synchronize()
    score_ms = (time.perf_counter() - t0) * 1e3
    idx = sorted(random.Random(1).sample(range(G), 16))
    worst = {"iou": 0.0, "gt_iou": 0.0, "far_noise_ratio": 0.0, "gauss_recall": 0.0}
    t0 = time.perf_counter()
    for i in idx:
        ref = R.compute_metrics(gt[i] < 128, R.binarise(full[i].cpu().numpy()), sigma=2.0)
        for k in worst:
            worst[k] = max(worst[k], abs(ref[k] - metrics[i][k]))
    oracle_ms_per_image = (time.perf_counter() - t0) * 1e3 / len(idx)
    block["eval_iou_parity"] = {
        "n_scored": G, "score_ms_incl_d2h": score_ms, "n_checked_against_oracle": len(idx), "indices": idx,
        "max_abs_diff": worst, "equal": all(v <= 1e-12 for v in worst.values()),
        "oracle_ms_per_image": oracle_ms_per_image,
        "mean": {k: float(np.mean([m[k] for m in metrics])) for k in worst},
        "ground_truth": "oracle.synth.render_gt: the condition's entity drawn 2 px wide on a white sheet "
                        "(random-init weights: the scores say nothing about image quality, only that the "
                        "device scorer and the reference's scorer agree on the gathered batch)"}
    return block


def step_parity(ctx, dev, unet_sd, y, vals, mask, coef, T, guidance):
    """Teacher-forced CFG step at t = T/2 on the TIMED batch, checked against the oracle port on four
    of its samples (checker only; outside every timed region)."""
    import torch
    from oracle import ref_torch as R
    B = y.shape[0]
    g = torch.Generator().manual_seed(7)
    x = torch.randn((B, 4, 28, 28), generator=g) * 3.0
    nz = torch.randn((B, 4, 28, 28), generator=g)
    tt = torch.full((B,), max(1, T // 2), dtype=torch.long)
    out, eps = ctx.denoise_cond_step(x.to(dev), tt.to(dev), y, 0, vals, mask, guidance, coef, noise=nz.to(dev),
                                     want_eps=True)
    pick = torch.tensor(sorted({0, B // 3, (2 * B) // 3, B - 1}))
    _, alphas, abars = R.schedule(T)
    with torch.no_grad():
        x_ref, eps_ref = R.denoise_cond(unet_sd, x[pick], tt[pick], y.cpu()[pick], guidance, 0, vals.cpu()[pick],
                                        mask.cpu()[pick], nz[pick], alphas, abars)
    eps, out = eps.cpu()[pick].double(), out.cpu()[pick].double()
    return {"t": int(tt[0]), "samples": pick.tolist(),
            "eps_rel_l2": float((eps - eps_ref.double()).norm() / eps_ref.double().norm()),
            "eps_max_abs_over_max": float((eps - eps_ref.double()).abs().max() / eps_ref.abs().max()),
            "x_max_abs_over_max": float((out - x_ref.double()).abs().max() / x_ref.abs().max()),
            "checker": "oracle port (torch CPU fp32) on 4 samples of the timed batch"}


def tf32_block(args, dev, unet, T, y, vals, mask, coef, peaks):
    """fp32-storage / TF32-operand mode on the benched batch: ms per denoise step inside the graph
    loop (200 steps), samples/s extrapolated, and the fraction of a TF32 peak measured here
    (cuBLAS fp32 matmul with TF32 allowed, 8192^3: best of 10 and sustained over ~1 s)."""
    import torch
    from dmb200 import context_for
    B = y.shape[0]
    ctx = context_for(dev, "tf32")
    ctx.ensure_unet(unet)
    n = min(T, 200)
    x = ctx.philox_normal(B, seed=3, sample_offset=0, t=0)
    ctx.sample_loop(x, y, 0, vals, mask, args.guidance, coef, T, 20, seed=5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ctx.sample_loop(x, y, 0, vals, mask, args.guidance, coef, T, n, seed=6)
    e1.record()
    torch.cuda.synchronize()
    ms_step = e0.elapsed_time(e1) / n
    # TF32 peak
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    N = 8192
    a = torch.randn((N, N), device=dev)
    b = torch.randn((N, N), device=dev)
    c = torch.empty((N, N), device=dev)
    for _ in range(3):
        torch.matmul(a, b, out=c)
    best = 1e9
    for _ in range(10):
        e0.record(); torch.matmul(a, b, out=c); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    reps = 60
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b, out=c)
    e1.record(); torch.cuda.synchronize()
    sustained = e0.elapsed_time(e1) / reps
    torch.backends.cuda.matmul.allow_tf32 = old
    del a, b, c
    fl = 2.0 * N ** 3
    peak_burst, peak_sus = fl / (best * 1e-3) / 1e12, fl / (sustained * 1e-3) / 1e12
    sps = B / (ms_step * 1e-3 * T)
    ach = sps * flop_per_sample(T) / 1e12
    return {"ms_per_denoise_step": ms_step, "samples_per_s_T_extrapolated": sps, "steps_timed": n, "batch": B,
            "whole_job_tflops": ach, "tf32_peak_tflops_burst": peak_burst, "tf32_peak_tflops_sustained": peak_sus,
            "whole_job_frac_of_sustained_tf32_peak": ach / peak_sus,
            "peak_how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), best of 10 / 60 back to back, CUDA events",
            "note": "decode excluded (loop only); bf16 is the headline mode"}


def batch1_eager(d, unet, dev, T, guidance, y1, v1, m1, n=200):
    """What generate_steps.py:180-189 does: Diffuser.denoise_cond called from Python once per step at
    batch 1 (the reference's range assertion syncs every step, as it does in the reference)."""
    import torch
    x = torch.randn((1, 4, 28, 28), device=dev)
    for i in range(5):
        t = torch.full((1,), T - i, device=dev, dtype=torch.long)
        x = d.denoise_cond(unet, x, t, y=y1, guidance_scale=guidance, null_label=0, cond_vals=v1, cond_mask=m1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        t = torch.full((1,), T - i, device=dev, dtype=torch.long)
        x = d.denoise_cond(unet, x, t, y=y1, guidance_scale=guidance, null_label=0, cond_vals=v1, cond_mask=m1)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3 / n


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
    B, T = args.batch, args.T
    # The CPU legs run FIRST, on rank 0, before this job holds a CUDA context or an NCCL communicator:
    # ranks that wait in an NCCL barrier poll with several host threads each, and an OpenMP workload
    # that shares cores with pollers loses a scheduler time slice at every parallel region (measured at
    # N = 2 with the legs at the end of the run: 4429 instead of 196 ms per step).  The other ranks wait
    # in the process group's rendezvous, which blocks on a socket.
    cpu = cpu_b1 = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_leg("baseline", batch=min(B, 16), n_steps=args.cpu_sample_steps, T=T, guidance=args.guidance)
        if not args.no_extras:
            # config 1 run outright on a 1-GPU launch (about a minute); bounded under torchrun
            cpu_b1 = cpu_leg("config1", T=T, guidance=args.guidance, n_steps=None if world == 1 else 20)
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(minutes=30))
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
    traffic, traffic_note = None, "no ncu capture of this build's GEMM kernel is committed"
    try:
        import hashlib
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            tj = json.load(f)
        h = hashlib.sha256()
        with open(os.path.join(PRODUCT, "csrc", "gemm_tc.cuh"), "rb") as f:   # the kernel's source
            h.update(f.read())
        # ... and the same launch schedule (tile selection lives in engine.cu): the GEMM launch count of a step
        if tj.get("source_sha16") == h.hexdigest()[:16] and tj.get("launches_per_step") == g_n // reps:
            traffic = tj.get("gemm_tc_kernel_dram_bytes_per_launch")
            traffic_note = tj.get("note", "")
        else:
            traffic_note = "profiles/roofline_traffic.json was captured on another build (kernel source hash or launch count differs): null"
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
        "traffic": traffic, "traffic_note": traffic_note,
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

        if not args.no_extras:
            batch1["eager_ms_per_step"] = batch1_eager(d, unet, dev, T, args.guidance, y1, v1, m1)
            batch1["eager_note"] = ("Diffuser.denoise_cond called from Python once per step (generate_steps.py:180-189), "
                                    "wall clock over 200 steps, includes the reference's per-step range assertion sync")

    parity = tf32 = None
    if rank == 0:
        unet_sd = {k: v.detach().cpu().clone() for k, v in unet.state_dict().items()}
        parity = step_parity(ctx, dev, unet_sd, y, vals, mask, coef, T, args.guidance)
        if not args.no_extras:
            tf32 = tf32_block(args, dev, unet, T, y, vals, mask, coef, peaks)

    # ---- extension N1: the strided deterministic schedule, 10 steps + decode, where the
    # VAE decoder is no longer negligible (device-resident inputs, like `value`)
    strided = None
    if rank == 0 and not args.no_extras and T % 10 == 0:
        coef10, stride10 = d.strided_coefficients(10)
        def strided_pass(i):
            x = ctx.philox_normal(B, seed=50 + i, sample_offset=0, t=0)
            ctx.sample_loop(x, y, 0, vals, mask, args.guidance, coef10, T, 10, seed=60 + i, t_stride=stride10)
            return ctx.vae_decode(x, want_f32=False, want_u8=True)[1]
        for i in range(3):
            strided_pass(i)
        torch.cuda.synchronize()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for i in range(10):
            strided_pass(10 + i)
        s1.record()
        torch.cuda.synchronize()
        ms10 = s0.elapsed_time(s1) / 10
        z = ctx.philox_normal(B, seed=1, sample_offset=0, t=0)
        ctx.vae_decode(z, want_f32=False, want_u8=True)
        torch.cuda.synchronize()
        s0.record()
        for _ in range(5):
            ctx.vae_decode(z, want_f32=False, want_u8=True)
        s1.record()
        torch.cuda.synchronize()
        dec_ms = s0.elapsed_time(s1) / 5
        strided = {"sample_steps": 10, "batch": B, "ms_per_batch": ms10, "samples_per_s": B / (ms10 * 1e-3),
                   "vae_decode_ms_per_batch": dec_ms, "decode_share": dec_ms / ms10,
                   "note": "dmb_sample_loop_strided (DDIM eta = 0, 10 of 1000 timesteps) + VAE decode to u8; "
                           "an opt-in extension, not the reference's schedule"}

    config4 = None
    if args.config4_batch > 0:
        config4 = run_config4(args, ctx, dev, world, rank, T, coef, barrier)

    if cpu_b1 is not None and batch1 is not None:
        batch1["cpu_ms_per_denoise_step"] = cpu_b1["ms_per_denoise_step"]

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
            "cpu_baseline_b1": cpu_b1,
            "ms_per_denoise_step": ms_total / args.steps / T,
            "batch1": batch1,
            "parity": parity,
            "tf32": tf32,
            "strided10": strided,
            "config4": config4,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.cpu_leg:
        run_cpu_leg(args.cpu_leg)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
