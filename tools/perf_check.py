#!/usr/bin/env python
"""Performance-regression gate (the reference's CI compares op benchmarks against a stored baseline,
tools/check_op_benchmark_result.py).  Compares one or more fresh bench.py JSON lines with the committed record under
profiles/ for the same (n_gpus, config) and fails when a fresh number is more than --tol worse.

    python bench.py --steps 6 --warmup 3 | tee gpurun_out/bench_now.json
    python tools/perf_check.py gpurun_out/bench_now.json --baseline profiles/bench_n1.json --tol 0.05
"""
import argparse
import json
import sys


def _load(path):
    rows = []
    with open(path) as fh:
        for line in fh:
            line = line.strip()
            if line.startswith("{"):
                try:
                    rows.append(json.loads(line))
                except json.JSONDecodeError:
                    pass
    if not rows:
        with open(path) as fh:
            d = json.load(fh)
        rows = d if isinstance(d, list) else [d]
    return rows


def _key(r):
    c = r.get("config", {})
    return (r.get("n_gpus"), c.get("model"), c.get("global_batch"), c.get("seq_len"), c.get("parallelism"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("fresh", nargs="+")
    ap.add_argument("--baseline", action="append", required=True)
    ap.add_argument("--tol", type=float, default=0.05, help="allowed relative slowdown")
    a = ap.parse_args()
    base = {}
    for b in a.baseline:
        for r in _load(b):
            base[_key(r)] = r
    bad = 0
    for f in a.fresh:
        for r in _load(f):
            ref = base.get(_key(r))
            if ref is None:
                print(f"[perf] no baseline for {_key(r)} ({f}): skipped")
                continue
            hib = r.get("higher_is_better", True)
            new, old = float(r["value"]), float(ref["value"])
            rel = (new - old) / old if hib else (old - new) / old
            status = "OK" if rel >= -a.tol else "REGRESSION"
            clocks = (r.get("clocks") or {}).get("sm_mhz"), (ref.get("clocks") or {}).get("sm_mhz")
            print(f"[perf] {status}: {_key(r)} {r.get('unit', '')} new={new:.1f} baseline={old:.1f} ({rel:+.1%}); "
                  f"sm_mhz new/baseline = {clocks[0]}/{clocks[1]}")
            bad += status != "OK"
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
