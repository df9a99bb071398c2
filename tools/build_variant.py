"""Build a variant of the kernel library with extra -D flags for one source file (A/B timing experiments on the GPU box):
    python tools/build_variant.py flash_attn.cu poly2 -DPDB_FA_POLY=2
writes build/variants/libpdb200_kernels_poly2.so; run with PDB200_KERNELS_LIB=build/variants/libpdb200_kernels_poly2.so."""
import subprocess, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from paddle_b200.lib import build as B

src_name, tag, *defs = sys.argv[1:]
B.build_kernels()
out_dir = B.ROOT / "build" / "variants"
out_dir.mkdir(parents=True, exist_ok=True)
objs = []
for src in sorted((B.ROOT / "csrc" / "kernels").glob("*.cu")):
    if src.name == src_name:
        obj = out_dir / f"{src.stem}_{tag}.o"
        subprocess.run([B.NVCC, *B.NVCC_FLAGS, *defs, "-c", str(src), "-o", str(obj)], check=True)
        objs.append(obj)
    else:
        objs.append(B._compile(src, B.NVCC, B.NVCC_FLAGS, False))
out = out_dir / f"libpdb200_kernels_{tag}.so"
subprocess.run([B.NVCC, "-shared", "-o", str(out), *map(str, objs), "-Xcompiler", "-fPIC", "-gencode", "arch=compute_100a,code=sm_100a"], check=True)
print(out)
