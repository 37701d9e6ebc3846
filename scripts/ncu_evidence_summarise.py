"""Turn the CSVs scripts/ncu_evidence.sh brought back into profiles/<tag>_*.md and profiles/roofline_traffic.json.
usage: python scripts/ncu_evidence_summarise.py <tag>      (e.g. r2_v12)
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
         "attention_mma_kernel": "attention_mma_kernel: the attention launches of one CFG step at batch 256 (L = 784, 196, 196, 49, 49 -- L = 9 runs on CUDA cores).",
         "attn_tail": "attn_tail kernels (out-projection + residual + LayerNorm + MLP + residual in one launch; C = 64/128 and the C = 256 form)."}
counts = {}
for k, note in notes.items():
    counts[k] = second_half_raw(f"{G}/ev_full_{k}.raw.csv", f"{G}/ev_full2_{k}.raw.csv")
    S.full(f"{G}/ev_full2_{k}.raw.csv", f"{P}/{tag}_ncu_full_{k.replace('_kernel', '')}.md", note)
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
