"""Turn the CSVs scripts/ncu_evidence.sh brought back into profiles/<tag>_*.md and profiles/roofline_traffic.json.
usage: python scripts/ncu_evidence_summarise.py <tag> [kernel ...]     (e.g. r2_v13 gemm_tc_kernel)
one_step.py runs two eager steps: every summary keeps the SECOND step's launches (the first builds plans)."""
import csv, hashlib, io, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import summarise_ncu_csv as S

tag = sys.argv[1]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def second_half_long(src, dst):
    """keep the second step of a long-format --metrics log: from the last emb_kernel launch (the first
    kernel of a step) to the last cfg_step_kernel launch (its last)"""
    txt = open(src).read()
    head, body = txt[:txt.index('"ID"')], txt[txt.index('"ID"'):]
    rows = list(csv.reader(io.StringIO(body)))
    name = rows[0].index("Kernel Name")
    first = max(int(r[0]) for r in rows[1:] if "emb_kernel" in r[name])
    last = max(int(r[0]) for r in rows[1:] if "cfg_step_kernel" in r[name])
    out = io.StringIO(); w = csv.writer(out, quoting=csv.QUOTE_ALL)
    w.writerow(rows[0])
    for r in rows[1:]:
        if first <= int(r[0]) <= last: w.writerow(r)
    open(dst, "w").write(head + out.getvalue())
    return last - first + 1


def second_half_raw(src, dst):
    rows = list(csv.reader(open(src)))
    data = rows[2:]
    assert len(data) % 2 == 0, f"{src}: {len(data)} launches"
    with open(dst, "w", newline="") as f:
        csv.writer(f, quoting=csv.QUOTE_ALL).writerows(rows[:2] + data[len(data) // 2:])
    return len(data) // 2


n = second_half_long(f"{G}/ev_step_metrics.csv", f"{G}/ev_step2.csv")
S.step(f"{G}/ev_step2.csv", f"{P}/{tag}_ncu_step_launches.md",
       f"DMB200_EAGER_GRAPH=0 python scripts/one_step.py 256 2  (second step: {n} launches)")
notes = {"gemm_tc_kernel": "gemm_tc_kernel: every launch of one CFG step at batch 256 (3x3 convs incl. kRow3 and pixel-major tiles, qkv projections).",
         "gn1_apply_kernel": "gn1_apply_kernel: every launch of one CFG step at batch 256 (statistics arrive as two 64-bit fixed-point sums per sample).",
         "attention_mma_kernel": "attention_mma_kernel: the attention launches of one CFG step at batch 256 (L = 784, 196, 196, 49, 49, 9).",
         "attn_tail": "attn_tail kernels (out-projection + residual + LayerNorm + MLP + residual in one launch; C = 64/128 and the C = 256 form).",
         "ln_qkv_kernel": "ln_qkv_kernel: LayerNorm + in-projection of the attention blocks with C = 64 / 128 at batch 256 (sa1, sa4, sa5, sa6).",
         "upcat_kernel": "upcat_kernel: skip copy + bilinear x2 + concat of the three Up stages (copy and interpolation in separate blocks).",
         "cfg_step_kernel": "cfg_step_kernel: out conv on both guidance halves + CFG combine + DDPM update (feature rows staged through shared memory).",
         "conv_in_mma_kernel": "conv_in_mma_kernel: 4 -> 64 input conv on mma.sync (all global loads of a CTA issued up front)."}
only = sys.argv[2:]    # optional: summarise only these kernels (the others keep their earlier captures)
counts = {}
for k, note in notes.items():
    if (only and k not in only) or not os.path.exists(f"{G}/ev_full_{k}.raw.csv"):
        continue
    counts[k] = second_half_raw(f"{G}/ev_full_{k}.raw.csv", f"{G}/ev_full2_{k}.raw.csv")
    S.full(f"{G}/ev_full2_{k}.raw.csv", f"{P}/{tag}_ncu_full_{k.replace('_kernel', '')}.md", note)
if not only or "score_masks" in only:
    S.full(f"{G}/ev_full_score_masks.raw.csv", f"{P}/{tag}_ncu_full_score_masks.md",
           "score_masks_kernel (dmb_score_masks, 4096 images of 224x224; scripts/bench_scoring.py).")

rows = list(csv.reader(open(f"{G}/ev_full2_gemm_tc_kernel.raw.csv")))
hdr, units, data = rows[0], rows[1], rows[2:]
def col(name, d):
    v = float(d[hdr.index(name)].replace(",", "")); u = units[hdr.index(name)]
    return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3}.get(u, 1.0)
dram = [col("dram__bytes_read.sum", d) + col("dram__bytes_write.sum", d) for d in data]
tus = [col("gpu__time_duration.sum", d) for d in data]
tp = [col("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", d) for d in data]
sha = hashlib.sha256(open(os.path.join(ROOT, "diffusion-model_b200", "csrc", "gemm_tc.cuh"), "rb").read()).hexdigest()[:16]
json.dump({"gemm_tc_kernel_dram_bytes_per_launch": sum(dram) / len(dram), "launches_per_step": len(dram),
           "dram_bytes_per_step": sum(dram),
           "time_weighted_tensor_pipe_active_pct": sum(a * b for a, b in zip(tp, tus)) / sum(tus),
           "ncu_time_us_per_step": sum(tus), "source_sha16": sha,
           "note": f"ncu --set full --clock-control none over the {len(dram)} gemm_tc_kernel launches of one CFG denoise step at batch 256 "
                   f"(profiles/{tag}_ncu_full_gemm_tc.md); dram__bytes_read.sum + dram__bytes_write.sum averaged per launch; valid while "
                   "csrc/gemm_tc.cuh hashes to source_sha16 and a step still has launches_per_step GEMM launches"},
          open(f"{P}/roofline_traffic.json", "w"), indent=1)
print(n, counts, open(f"{P}/roofline_traffic.json").read())


def bench_launches(src, dst):
    """launch list of the bench command itself (gpu__time_duration.sum only): launches, total time, share per kernel"""
    import collections, re
    txt = open(src).read()
    rows = list(csv.DictReader(io.StringIO(txt[txt.index('"ID"'):])))
    agg, tot = collections.OrderedDict(), 0.0
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", "")) * {"us": 1, "ms": 1e3, "ns": 1e-3}.get(r["Metric Unit"], 1)
        n = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("dmb::", "")
        a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += v; tot += v
    ours = sum(v[1] for k, v in agg.items() if not k.startswith("at::") and "nccl" not in k and "cub::" not in k)
    with open(dst, "w") as f:
        f.write("# ncu launch list of the bench command (`--metrics gpu__time_duration.sum --clock-control none`)\n\n"
                "command: `python bench.py --steps 1 --warmup 3 --T 4 --no-cpu-baseline --no-extras` (after the same command exited 0 "
                "without ncu; first 3000 launches).  Times are cold-cache and serialised: compare SHARES with the bench line's kernel "
                "groups, not absolutes.\n\n"
                f"{sum(v[0] for v in agg.values())} launches, {tot / 1e3:.1f} ms total; the repo's own kernels: {100 * ours / tot:.1f} % of the time\n\n"
                "| kernel | launches | total us | share |\n|---|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k[:100]}` | {v[0]} | {v[1]:.1f} | {100 * v[1] / tot:.1f}% |\n")


if os.path.exists(f"{G}/ev_bench_launches.csv"):
    bench_launches(f"{G}/ev_bench_launches.csv", f"{P}/{tag}_ncu_bench_launches.md")
