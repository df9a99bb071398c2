"""In-tree build of the native libraries (no GPU needed: nvcc cross-compiles sm_100a).

* ``libpdb200_kernels.so``  <- csrc/kernels/*.cu   (hand-written sm_100a kernels, C ABI, include/pdb200_kernels.h)
* ``libpdb200_runtime.so``  <- csrc/runtime/*.cc   (TCP store, comm watchdog, reducer bucketing, symmetric-memory
                                                    arena; C ABI, include/pdb200_runtime.h)

Objects are cached under build/ keyed by source mtime + flags; files compile in parallel.
Run ``python -m paddle_b200.lib.build`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
LIB_DIR = Path(__file__).resolve().parent
BUILD_DIR = ROOT / "build"
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
NVCC = os.path.join(CUDA_HOME, "bin", "nvcc")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
              "-I" + str(ROOT / "include"), "-I" + str(ROOT / "csrc" / "kernels")]
CXX_FLAGS = ["-O2", "-g", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-pthread", "-I" + str(ROOT / "include"),
             "-I" + os.path.join(CUDA_HOME, "include")]


def _stamp(src: Path, flags) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for inc in sorted(list((ROOT / "csrc" / "kernels").glob("*.cuh")) + list((ROOT / "include").glob("*.h")) +
                      list((ROOT / "csrc" / "runtime").glob("*.h"))):
        h.update(inc.read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()[:16]


def _compile(src: Path, compiler, flags, verbose) -> Path:
    BUILD_DIR.mkdir(exist_ok=True)
    obj = BUILD_DIR / (src.stem + "." + _stamp(src, flags) + ".o")
    if obj.exists():
        return obj
    for old in BUILD_DIR.glob(src.stem + ".*.o"):
        old.unlink()
    cmd = [compiler, *flags, "-c", str(src), "-o", str(obj)]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {src}\n{r.stdout[-4000:]}\n{r.stderr[-8000:]}")
    return obj


def _link(objs, out: Path, extra, compiler) -> None:
    key = hashlib.sha1(" ".join(sorted(o.name for o in objs)).encode()).hexdigest()[:16]
    tag = out.with_suffix(".tag")
    if out.exists() and tag.exists() and tag.read_text() == key:
        return
    cmd = [compiler, "-shared", "-o", str(out), *[str(o) for o in objs], *extra]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed: {out}\n{r.stderr[-8000:]}")
    tag.write_text(key)


def build_kernels(verbose=False) -> Path:
    srcs = sorted((ROOT / "csrc" / "kernels").glob("*.cu"))
    with ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(lambda s: _compile(s, NVCC, NVCC_FLAGS, verbose), srcs))
    out = LIB_DIR / "libpdb200_kernels.so"
    _link(objs, out, ["-Xcompiler", "-fPIC", "-gencode", "arch=compute_100a,code=sm_100a"], NVCC)
    return out


def build_runtime(verbose=False) -> Path | None:
    srcs = sorted((ROOT / "csrc" / "runtime").glob("*.cc"))
    if not srcs:
        return None
    with ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(lambda s: _compile(s, "g++", CXX_FLAGS, verbose), srcs))
    out = LIB_DIR / "libpdb200_runtime.so"
    _link(objs, out, ["-pthread", "-L" + os.path.join(CUDA_HOME, "lib64"), "-lcudart", "-ldl"], "g++")
    return out


def build_all(verbose=False):
    k = build_kernels(verbose)
    r = build_runtime(verbose)
    return k, r


if __name__ == "__main__":
    k, r = build_all(verbose="-v" in sys.argv)
    print("built", k, r)
