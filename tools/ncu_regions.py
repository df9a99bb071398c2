"""Warp-stall samples of one kernel aggregated over consecutive SASS regions (which phase of the kernel waits, and why).
usage: ncu -i rep --page source --csv --kernel-name regex:NAME > src.csv; python tools/ncu_regions.py src.csv [chunk]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 50
# the export may hold several kernels: keep the first section
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
end = starts[1] if len(starts) > 1 else len(rows)
hdr = rows[starts[0] + 1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[starts[0] + 2:end] if len(r) == len(hdr)]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
print(rows[starts[0]][1][:100], "total samples", tot)
for a in range(0, len(data), chunk):
    seg = data[a:a + chunk]
    n = sum(int(r[ix["# Samples"]] or 0) for r in seg)
    if n / max(tot, 1) < 0.012:
        continue
    ex = sum(int(r[ix["Instructions Executed"]] or 0) for r in seg)
    why = sorted(((sum(int(r[ix[s]] or 0) for r in seg), s[6:]) for s in stalls), reverse=True)[:3]
    ops = {}
    for r in seg:
        sp = r[ix["Source"]].split()
        op = sp[0] if sp else ""
        if op.startswith("@") and len(sp) > 1:
            op = sp[1]
        ops[op] = ops.get(op, 0) + 1
    top = sorted(ops.items(), key=lambda kv: -kv[1])[:4]
    print(f"{a:5d} {100 * n / tot:5.1f}% exec={ex:9d} {why} {top}")
