"""Summaries of ncu data that was converted to CSV on the GPU box (the .ncu-rep files are too large
to bring back):
  python scripts/summarise_ncu_csv.py full  gpurun_out/prof_x.raw.csv profiles/<name>.md ["note"]
      one table per captured launch from `ncu -i prof.ncu-rep --page raw --csv`
  python scripts/summarise_ncu_csv.py step  gpurun_out/step_metrics.csv profiles/<name>.md "<command>"
      every launch of one step from `ncu --metrics ... --csv --log-file` (long format)
"""
import collections, csv, io, re, sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__inst_executed.sum", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__cycles_active.avg"]


def short(name):
    return re.sub(r"\(.*", "", name).replace("void ", "").replace("dmb::", "")


def full(src, dst, note=""):
    rows = list(csv.reader(open(src)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full --clock-control none: {src.split('/')[-1]}\n\n{note}\n\n")
        for d in data:
            f.write(f"## `{short(d[hdr.index('Kernel Name')])[:110]}`  grid {d[hdr.index('Grid Size')]} block {d[hdr.index('Block Size')]}\n\n"
                    "| metric | value | unit |\n|---|---:|---|\n")
            for m in KEEP:
                if m in hdr:
                    f.write(f"| {m} | {d[hdr.index(m)]} | {units[hdr.index(m)]} |\n")
            st = sorted(((float(d[hdr.index(h)] or 0), h.split("stalled_")[1].split("_per_issue")[0]) for h in stalls), reverse=True)
            f.write("| warp stalls per issue (top 6) | " + ", ".join(f"{n} {v:.2f}" for v, n in st[:6]) + " | |\n\n")


def step(src, dst, cmd):
    txt = open(src).read()
    rows = list(csv.DictReader(io.StringIO(txt[txt.index('"ID"'):])))
    launches = collections.OrderedDict()
    for r in rows:
        L = launches.setdefault(r["ID"], {"name": short(r["Kernel Name"]), "grid": r["Grid Size"], "block": r["Block Size"]})
        L[r["Metric Name"]] = float(r["Metric Value"].replace(",", "")) * {"us": 1e3, "ms": 1e6, "ns": 1.0, "Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(r["Metric Unit"], 1.0)
    tot = sum(L["gpu__time_duration.sum"] for L in launches.values())
    agg = collections.OrderedDict()
    for L in launches.values():
        a = agg.setdefault(L["name"], [0, 0.0, 0.0])
        a[0] += 1; a[1] += L["gpu__time_duration.sum"]; a[2] += L["dram__bytes_read.sum"] + L["dram__bytes_write.sum"]
    with open(dst, "w") as f:
        f.write(f"# ncu launch list of one CFG denoise step (--clock-control none; cold-cache, serialised: compare SHARES)\n\n"
                f"command: `{cmd}`\n\n{len(launches)} launches, {tot / 1e3:.1f} us total\n\n"
                "| kernel | launches | total us | share | DRAM MB (read + write) |\n|---|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k[:90]}` | {v[0]} | {v[1] / 1e3:.1f} | {100 * v[1] / tot:.1f}% | {v[2] / 1e6:.1f} |\n")
        f.write("\n## every launch\n\n| # | kernel | grid | block | us | DRAM read MB | DRAM write MB | DRAM % of peak | tensor pipe % | issue % | xbar->L1 MB |\n"
                "|---|---|---|---|---:|---:|---:|---:|---:|---:|---:|\n")
        for i, L in launches.items():
            f.write(f"| {i} | `{L['name'][:70]}` | {L['grid']} | {L['block']} | {L['gpu__time_duration.sum'] / 1e3:.1f} | "
                    f"{L['dram__bytes_read.sum'] / 1e6:.2f} | {L['dram__bytes_write.sum'] / 1e6:.2f} | "
                    f"{L.get('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 0):.1f} | "
                    f"{L.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 0):.1f} | "
                    f"{L.get('sm__inst_issued.avg.pct_of_peak_sustained_active', 0):.1f} | "
                    f"{L.get('l1tex__m_xbar2l1tex_read_bytes.sum', 0) / 1e6:.1f} |\n")
    return launches


if __name__ == "__main__":
    if sys.argv[1] == "full":
        full(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        step(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
