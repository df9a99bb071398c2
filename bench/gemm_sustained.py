"""Sustained (power-limited) GEMM throughput: each variant runs back to back for ~1.5 s, the last second is timed with
CUDA events, and nvidia-smi samples SM clock and board power meanwhile.  Compares cuBLAS (torch.matmul) with the
CTA-pair tcgen05 kernel at several raster band heights.  One JSON line per (shape, variant)."""
import json
import os
import subprocess
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200 import lib  # noqa: E402
from paddle_b200.ops import gemm as G  # noqa: E402


class Smi:
    def __init__(self):
        self.rows = []
        self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100", "-i", "0"],
                                  stdout=subprocess.PIPE, text=True)
        threading.Thread(target=self._rd, daemon=True).start()

    def _rd(self):
        for l in self.p.stdout:
            try:
                a, b = l.split(",")
                self.rows.append((time.time(), float(a), float(b)))
            except ValueError:
                pass

    def window(self, t0, t1):
        r = [(c, w) for t, c, w in self.rows if t0 <= t <= t1]
        if not r:
            return None, None
        r.sort()
        return r[len(r) // 2][0], round(sum(w for _, w in r) / len(r), 1)


def sustained(fn, secs=1.0, warm=0.6):
    t_end = time.time() + warm
    while time.time() < t_end:
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    t0 = time.time()
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, t0, time.time()


def main():
    smi = Smi()
    T = int(os.environ.get("GEMM_TOKENS", "8192"))
    shapes = [("qkv fwd", T, 12288, 4096, False, False), ("gate_up fwd", T, 22016, 4096, False, False),
              ("down fwd", T, 4096, 11008, False, False), ("gate_up dgrad", T, 4096, 22016, False, True),
              ("gate_up wgrad", 4096, 22016, T, True, False), ("down wgrad", 11008, 4096, T, True, False)]
    variants = os.environ.get("GEMM_VARIANTS", "cublas,g16,m8,m16").split(",")
    for name, M, N, K, ta, tb in shapes:
        nbuf = 3
        As = [torch.randn((K, M) if ta else (M, K), device="cuda").bfloat16() for _ in range(nbuf)]
        Bs = [torch.randn((N, K) if tb else (K, N), device="cuda").bfloat16() for _ in range(nbuf)]
        out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        i = [0]
        for v in variants:
            if v == "cublas":
                def fn():
                    i[0] = (i[0] + 1) % nbuf
                    a = As[i[0]].t() if ta else As[i[0]]
                    b = Bs[i[0]].t() if tb else Bs[i[0]]
                    torch.matmul(a, b, out=out)
            else:
                from paddle_b200.framework import flags
                if v.startswith("g"):
                    flags.set_flags({"gemm_kernel": "pair"})
                    lib.kernels.pdb200_gemm_set_group_m(int(v[1:]))
                else:                                   # "m16": macro-tile kernel, band 16
                    flags.set_flags({"gemm_kernel": "macro"})
                    lib.kernels.pdb200_gemm_set_band(int(v[1:]))

                def fn():
                    i[0] = (i[0] + 1) % nbuf
                    G.gemm(As[i[0]], Bs[i[0]], ta, tb, out=out)
            ms, t0, t1 = sustained(fn)
            clk, pw = smi.window(t0 + 0.2, t1)
            print(json.dumps(dict(shape=name, M=M, N=N, K=K, variant=v, ms=round(ms, 4), tflops=round(2.0 * M * N * K / ms / 1e9, 1),
                                  sm_mhz=clk, watts=pw)), flush=True)
    smi.p.terminate()


if __name__ == "__main__":
    main()
