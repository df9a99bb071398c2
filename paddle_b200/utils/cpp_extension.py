"""paddle.utils.cpp_extension (reference: python/paddle/utils/cpp_extension/{cpp_extension,extension_utils}.py): build
user C++ / CUDA operators against this framework.

Custom operators here are what they are in the rest of the tree: functions over ``torch.Tensor`` in a shared library,
built with the torch extension tool chain.  ``CUDAExtension`` and ``load`` always compile for sm_100a (the one target
of this framework; ``-gencode arch=compute_100a,code=sm_100a`` is appended unless the caller names an arch) and put
this framework's C-ABI headers (``paddle.sysconfig.get_include()``) on the include path."""
from __future__ import annotations

import os

__all__ = ["CppExtension", "CUDAExtension", "load", "setup", "get_build_directory"]

_SM100A = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo"]


def _with_arch(flags):
    flags = list(flags or [])
    return flags if any("arch=" in f or f.startswith("-arch") for f in flags) else flags + _SM100A


def _include_dirs(extra):
    from .. import sysconfig
    return list(extra or []) + [sysconfig.get_include()]


def get_build_directory(verbose=False):
    """Where JIT-built extensions go: ``$PADDLE_EXTENSION_DIR`` or ``~/.cache/paddle_extensions``."""
    root = os.environ.get("PADDLE_EXTENSION_DIR") or os.path.join(os.path.expanduser("~"), ".cache", "paddle_extensions")
    os.makedirs(root, exist_ok=True)
    if verbose:
        print(f"$PADDLE_EXTENSION_DIR is not set, using path: {root} by default." if "PADDLE_EXTENSION_DIR" not in os.environ
              else f"using $PADDLE_EXTENSION_DIR: {root}")
    return root


def CppExtension(sources, *args, **kwargs):  # noqa: N802
    """A setuptools Extension for C++ sources (``name`` is taken from ``setup(name=...)`` when omitted)."""
    from torch.utils import cpp_extension as T
    name = kwargs.pop("name", None) or "_paddle_cpp_extension"
    kwargs["include_dirs"] = _include_dirs(kwargs.get("include_dirs"))
    return T.CppExtension(name, list(sources), *args, **kwargs)


def CUDAExtension(sources, *args, **kwargs):  # noqa: N802
    from torch.utils import cpp_extension as T
    name = kwargs.pop("name", None) or "_paddle_cuda_extension"
    kwargs["include_dirs"] = _include_dirs(kwargs.get("include_dirs"))
    compile_args = kwargs.get("extra_compile_args") or {}
    if isinstance(compile_args, dict):
        compile_args = dict(compile_args)
        compile_args["nvcc"] = _with_arch(compile_args.get("nvcc"))
        compile_args.setdefault("cxx", [])
    kwargs["extra_compile_args"] = compile_args
    return T.CUDAExtension(name, list(sources), *args, **kwargs)


def setup(**attr):
    """``setuptools.setup`` with the torch build-extension command; ``ext_modules`` may be one Extension or a list, and
    extensions created without a name take ``name``."""
    import setuptools
    from torch.utils import cpp_extension as T
    name = attr.get("name")
    if name is None:
        raise ValueError("setup() needs name=")
    ext = attr.get("ext_modules")
    if ext is None:
        raise ValueError("setup() needs ext_modules=")
    ext = list(ext) if isinstance(ext, (list, tuple)) else [ext]
    for e in ext:
        if e.name in ("_paddle_cpp_extension", "_paddle_cuda_extension"):
            e.name = name
    attr["ext_modules"] = ext
    cmdclass = dict(attr.get("cmdclass") or {})
    cmdclass.setdefault("build_ext", T.BuildExtension.with_options(no_python_abi_suffix=True))
    attr["cmdclass"] = cmdclass
    return setuptools.setup(**attr)


def load(name, sources, extra_cxx_cflags=None, extra_cuda_cflags=None, extra_ldflags=None, extra_include_paths=None,
         extra_library_paths=None, build_directory=None, verbose=False):
    """Compile ``sources`` just in time and import the module."""
    from torch.utils import cpp_extension as T
    build_directory = build_directory or os.path.join(get_build_directory(), name)
    os.makedirs(build_directory, exist_ok=True)
    has_cuda = any(str(s).endswith((".cu", ".cuh")) for s in sources)
    ldflags = list(extra_ldflags or []) + [f"-L{p}" for p in (extra_library_paths or [])]
    return T.load(name=name, sources=list(sources), extra_cflags=list(extra_cxx_cflags or []),
                  extra_cuda_cflags=_with_arch(extra_cuda_cflags) if has_cuda else None, extra_ldflags=ldflags,
                  extra_include_paths=_include_dirs(extra_include_paths), build_directory=build_directory, verbose=verbose,
                  with_cuda=has_cuda or None)
