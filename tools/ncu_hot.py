"""Top stalled SASS instructions of one kernel from `ncu -i rep --page source --csv --kernel-name regex:...` output."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(int(r[ix[s]] or 0) for r in data) for s in stalls}
print("total samples", tot)
print("by reason:", sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
top = sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for r in top:
    n = int(r[ix["# Samples"]] or 0)
    why = sorted(((int(r[ix[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"{100*n/tot:5.1f}%  {r[ix['Source']].strip()[:70]:70s} {why}")
