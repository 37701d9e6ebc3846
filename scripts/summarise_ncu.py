"""Turns the ncu outputs a gpurun visit brought back (gpurun_out/) into the small tracked
summaries under profiles/.
  python scripts/summarise_ncu.py launches gpurun_out/launches.csv profiles/<name>.md "<command>"
  python scripts/summarise_ncu.py full gpurun_out/prof_x.ncu-rep profiles/<name>.md
"""
import collections, csv, io, re, subprocess, sys

KEEP = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
        "smsp__cycles_active.avg", "sm__inst_executed_pipe_tensor_subpipe_hmma.sum",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]


def launches(src, dst, cmd):
    txt = open(src).read()
    rows = list(csv.DictReader(io.StringIO(txt[txt.index('"ID"'):])))
    agg = collections.OrderedDict()
    for r in rows:
        k = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r["Metric Value"])
    tot = sum(v[1] for v in agg.values())
    with open(dst, "w") as f:
        f.write(f"# ncu launch list (gpu__time_duration.sum, --clock-control none)\n\ncommand: `{cmd}`\n\n"
                f"{len(rows)} launches, {tot / 1e3:.1f} us total (cold-cache, serialised: compare shares)\n\n"
                "| kernel | launches | total us | share | avg us |\n|---|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {v[0]} | {v[1] / 1e3:.1f} | {100 * v[1] / tot:.1f}% | {v[1] / v[0] / 1e3:.1f} |\n")
        f.write("\n## every launch\n\n| # | kernel | grid | block | ns |\n|---|---|---|---|---:|\n")
        for r in rows:
            k = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
            f.write(f"| {r['ID']} | `{k}` | {r['Grid Size']} | {r['Block Size']} | {float(r['Metric Value']):.0f} |\n")


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full capture: {src.split('/')[-1]}\n\n")
        for d in data:
            name = d[hdr.index("Kernel Name")]
            f.write(f"## `{name[:100]}`  grid {d[hdr.index('Grid Size')]} block {d[hdr.index('Block Size')]}\n\n"
                    "| metric | value | unit |\n|---|---:|---|\n")
            for m in KEEP:
                if m in hdr:
                    i = hdr.index(m)
                    f.write(f"| {m} | {d[i]} | {units[i]} |\n")
            f.write("\n")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        full(sys.argv[2], sys.argv[3])
