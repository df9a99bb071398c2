"""Count the Blackwell-specific SASS mnemonics per kernel of libpdb200_kernels.so (evidence that the tcgen05 / TMA /
NVLS paths are real): UTCHMMA(.2CTA) = tcgen05.mma, UTMALDG / UTMASTG = TMA load / store, LDTM/STTM = tcgen05.ld/st, UTCBAR = commit,
SYNCS = mbarrier, LDGMC = multimem.ld_reduce (NVLS in-switch reduction), *.STRONG.SYS = system-scope peer loads /
stores and flags, FFMA2 = packed fp32.  Kernels without any of these are omitted.
usage: python tools/sass_evidence.py > profiles/sass_evidence.md"""
import re, subprocess, sys
from collections import OrderedDict

so = sys.argv[1] if len(sys.argv) > 1 else "paddle_b200/lib/libpdb200_kernels.so"
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA.2CTA", "UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "SYNCS", "LDGMC", "STG.E.128.STRONG.SYS", "LDG.E.STRONG.SYS", "FFMA2", "MUFU.EX2"]
PROOF = KEYS[:11]
per = OrderedDict()
cur = None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        per[cur] = {k: 0 for k in KEYS}
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    for k in KEYS:
        if k == "UTCHMMA":
            if op.startswith("UTCHMMA") and ".2CTA" not in op:
                per[cur][k] += 1
        elif op.startswith(k):
            per[cur][k] += 1
dem = subprocess.run(["c++filt"], input="\n".join(per.keys()), capture_output=True, text=True).stdout.splitlines()
print("# SASS evidence (cuobjdump -sass %s)\n" % so)
print("| kernel | " + " | ".join(KEYS) + " |")
print("|---|" + "---:|" * len(KEYS))
for (name, cnt), d in zip(per.items(), dem):
    if not any(cnt[k] for k in PROOF if k != 'SYNCS'):
        continue
    short = re.sub(r"\(.*", "", d).replace("void ", "").replace("pdb::", "")[:70]
    print(f"| `{short}` | " + " | ".join(str(cnt[k]) if cnt[k] else "" for k in KEYS) + " |")
