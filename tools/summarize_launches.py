"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table (markdown).

usage: python tools/summarize_launches.py gpurun_out/launches.csv [skip_first_n_launches] > profiles/launch_summary.md
The launch list is the evidence of WHICH kernels run in a training step and how the step time splits between our
sm_100a kernels and library kernels; times under ncu are serialised per kernel, so they are a breakdown, not a bench."""
import csv
import re
import sys
from collections import defaultdict


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = re.sub(r"<.*", "", name)
    name = name.replace("void ", "")
    return name.strip()[:90]


def family(name):
    n = name.lower()
    if any(t in n for t in ("pdb::", "pdb200", "tc::", "tc2::", "fa::", "fab::", "nvl::", "gemm_tc", "flash_fwd", "flash_bwd")):
        return "ours (sm_100a kernels of this repo)"
    if "cutlass" in n or "cublas" in n or "nvjet" in n or "gemm" in n or "xmma" in n:
        return "library GEMM"
    if "cudnn" in n or "fmha" in n or "flash" in n or "sdpa" in n:
        return "library attention"
    if "nccl" in n:
        return "NCCL"
    return "torch elementwise/other"


def main():
    path = sys.argv[1]
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    per = defaultdict(lambda: [0, 0.0, ""])
    total = 0.0
    with open(path, newline="") as f:
        rows = [r for r in csv.reader(f) if len(r) > 14 and r[0].isdigit()]
    for r in rows:
        if int(r[0]) < skip:
            continue
        full, unit, val = r[4], r[13], float(r[14].replace(",", ""))
        ns = val * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9, "nsecond": 1, "usecond": 1e3, "msecond": 1e6, "second": 1e9}.get(unit, 1)
        k = short(full)
        per[k][0] += 1
        per[k][1] += ns
        per[k][2] = family(full)
        total += ns
    fam = defaultdict(float)
    for k, (c, ns, fm) in per.items():
        fam[fm] += ns
    print(f"# Kernel launch list summary ({path}, launches >= {skip})\n")
    print(f"total kernel time {total / 1e6:.2f} ms over {sum(v[0] for v in per.values())} launches\n")
    print("| family | ms | share |\n|---|---:|---:|")
    for fm, ns in sorted(fam.items(), key=lambda kv: -kv[1]):
        print(f"| {fm} | {ns / 1e6:.2f} | {100 * ns / total:.1f}% |")
    print("\n| kernel | family | launches | total ms | share |\n|---|---|---:|---:|---:|")
    for k, (c, ns, fm) in sorted(per.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"| `{k}` | {fm} | {c} | {ns / 1e6:.3f} | {100 * ns / total:.1f}% |")


if __name__ == "__main__":
    main()
