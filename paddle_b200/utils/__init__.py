"""paddle.utils (reference: python/paddle/utils/)."""
from . import unique_name  # noqa: F401


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
