"""paddle.utils (reference: python/paddle/utils/)."""
from . import cpp_extension, unique_name  # noqa: F401


def try_import(name):
    import importlib
    return importlib.import_module(name)


def flatten(nest):
    out = []

    def rec(x):
        if isinstance(x, (list, tuple)):
            for v in x:
                rec(v)
        elif isinstance(x, dict):
            for k in sorted(x):
                rec(x[k])
        else:
            out.append(x)
    rec(nest)
    return out


def map_structure(fn, *structs):
    s0 = structs[0]
    if isinstance(s0, (list, tuple)):
        return type(s0)(map_structure(fn, *xs) for xs in zip(*structs))
    if isinstance(s0, dict):
        return {k: map_structure(fn, *(s[k] for s in structs)) for k in s0}
    return fn(*structs)


class deprecated:  # decorator no-op
    def __init__(self, *a, **k):
        pass

    def __call__(self, fn):
        return fn


def run_check():
    """Install check (reference: python/paddle/utils/install_check.py): one tiny training step on the available
    device, through the native kernels when a GPU is present, and a report of what was used."""
    import torch

    import paddle_b200 as paddle
    dev = "gpu" if torch.cuda.is_available() else "cpu"
    print(f"Running verify PaddlePaddle(B200) program ... device: {dev}")
    lin = paddle.nn.Linear(8, 4)
    if dev == "gpu":
        lin.to("cuda")
    x = torch.randn(2, 8, device="cuda" if dev == "gpu" else "cpu")
    opt = paddle.optimizer.SGD(0.1, parameters=lin.parameters())
    lin(x).sum().backward()
    opt.step()
    if dev == "gpu":
        from paddle_b200 import lib
        print(f"native kernel library: {'loaded' if lib.have_kernels() else 'MISSING'}; GPUs visible: {torch.cuda.device_count()}")
    print("PaddlePaddle(B200) works well on 1 device.\nPaddlePaddle(B200) is installed successfully!")


def require_version(min_version, max_version=None):
    """The API level this framework tracks is the reference snapshot's (develop, 0.0.0): any requirement passes unless
    it is malformed."""
    for v in (min_version, max_version):
        if v is not None and (not isinstance(v, str) or not all(p.isdigit() for p in v.split(".") if p)):
            raise ValueError(f"version must look like '1.2.3', got {v!r}")
