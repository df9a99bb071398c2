"""Per-tile timeline of CTA 0 of the forward flash kernel (library built with -DPDB_FA_TRACE, see tools/build_variant.py)."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200 import lib, ops

so = ctypes.CDLL(os.environ["PDB200_KERNELS_LIB"])
b, s, h, d = 2, 4096, 32, 128
q, k, v = (torch.randn(b, s, h, d, device="cuda").bfloat16() for _ in range(3))
for _ in range(3):
    ops.flash_attention(q, k, v, causal=True)
buf = torch.zeros(3 * 4096, dtype=torch.int64, device="cuda")
so.pdb200_flash_attn_set_trace.argtypes = [ctypes.c_void_p]
assert so.pdb200_flash_attn_set_trace(buf.data_ptr()) == 0
ops.flash_attention(q, k, v, causal=True)
torch.cuda.synchronize()
so.pdb200_flash_attn_set_trace(None)
t = buf.cpu().view(3, 4096)
ev = []
for role in range(3):
    for x in t[role].tolist():
        if x:
            ev.append((x & ((1 << 56) - 1), role, (x >> 56) & 0xff))
ev.sort()
t0 = ev[0][0]
names = {0: "wait_S", 1: "S_ready", 2: "ld_done", 3: "math_done", 4: "o_ok", 5: "P_arrived", 6: "merge_read", 7: "item_done",
         9: "wait_P", 10: "P_seen", 11: "pv_issued", 12: "K_seen", 13: "qk_issued"}
out = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "attn_trace.txt"), "w")
for c, role, tag in ev[:1500]:
    out.write(f"{c - t0:8d} {'MMA' if role == 0 else 'grp' + 'AB'[role - 1]:5s} {names.get(tag, tag)}\n")
out.close()
print("events", len(ev))
