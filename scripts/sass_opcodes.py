"""Per-kernel counts of the SASS opcodes that show which hardware path a kernel uses
(cuobjdump -sass on the shipped library; runs without a GPU):
  UTCHMMA / UTCQMMA ...  tcgen05.mma          LDTM / STTM   tcgen05.ld / st (TMEM)
  UTMALDG / UTMASTG      TMA load / store     HMMA          mma.sync (legacy warp-level path)
  MUFU                   special function     LDSM          ldmatrix
    python scripts/sass_opcodes.py > profiles/r2_sass_opcodes.md
"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "diffusion-model_b200", "dmb200", "libdmb200.so")
COLS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "LDSM", "MUFU", "FFMA2", "ATOMG", "RED"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = counts.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
            cur["_total"] += 1
    names = demangle(list(counts))
    print("# SASS opcode counts per kernel (`cuobjdump -sass diffusion-model_b200/dmb200/libdmb200.so`, sm_100a)\n")
    print("Static instruction counts (not executed counts).  tcgen05.mma = UTC*MMA, TMEM access = LDTM / STTM, TMA = UTMALDG / "
          "UTMASTG, mbarrier = SYNCS, legacy warp-level MMA = HMMA.\n")
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print("Library totals: " + ", ".join(f"{k} {tot[k]}" for k in COLS if tot[k]) + f"; {len(counts)} kernels, {tot['_total']} instructions\n")
    print("| kernel | instr | " + " | ".join(COLS) + " |\n|---|---:|" + "---:|" * len(COLS))
    def key(kv):
        return (-(kv[1]["UTCHMMA"] + kv[1]["UTCQMMA"]), -kv[1]["HMMA"], -kv[1]["_total"])
    for mangled, c in sorted(counts.items(), key=key):
        n = re.sub(r"\(.*", "", names.get(mangled, mangled)).replace("void ", "").replace("dmb::", "")
        n = n.replace("__nv_bfloat16", "bf16")
        print(f"| `{n[:80]}` | {c['_total']} | " + " | ".join(str(c[k]) if c[k] else "" for k in COLS) + " |")


if __name__ == "__main__":
    main()
