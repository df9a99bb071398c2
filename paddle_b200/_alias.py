"""Import alias: ``paddle.*`` resolves to the very same module objects as ``paddle_b200.*``.

A user of the reference writes ``import paddle.distributed.fleet as fleet``; here that
statement must hand back ``paddle_b200.distributed.fleet`` itself (not a second copy of the
module, which would duplicate global state such as the default process group).  A meta-path
finder maps every ``paddle[.x.y]`` name to the already-imported ``paddle_b200[.x.y]`` module.
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.util
import sys

_REAL = "paddle_b200"
_ALIAS = "paddle"


def _real_name(fullname: str) -> str:
    return _REAL + fullname[len(_ALIAS):]


class _AliasLoader(importlib.abc.Loader):
    def create_module(self, spec):
        mod = importlib.import_module(_real_name(spec.name))
        # module_from_spec overwrites __spec__; remember the genuine one to restore it.
        self._saved = getattr(mod, "__spec__", None)
        return mod

    def exec_module(self, module):
        saved = getattr(self, "_saved", None)
        if saved is not None:
            try:
                module.__spec__ = saved
            except Exception:
                pass

    # runpy (`python -m paddle.distributed.launch`) asks the loader for code.
    def get_code(self, fullname):
        real = importlib.util.find_spec(_real_name(fullname))
        if real is None or real.loader is None:
            return None
        return real.loader.get_code(_real_name(fullname))

    def is_package(self, fullname):
        real = importlib.util.find_spec(_real_name(fullname))
        return bool(real and real.submodule_search_locations is not None)

    def get_filename(self, fullname):
        real = importlib.util.find_spec(_real_name(fullname))
        return real.origin if real else None


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname != _ALIAS and not fullname.startswith(_ALIAS + "."):
            return None
        real = _real_name(fullname)
        try:
            real_spec = importlib.util.find_spec(real)
        except (ImportError, ValueError):
            return None
        if real_spec is None:
            return None
        spec = importlib.util.spec_from_loader(
            fullname, _AliasLoader(), origin=real_spec.origin,
            is_package=real_spec.submodule_search_locations is not None)
        if real_spec.submodule_search_locations is not None:
            spec.submodule_search_locations = list(real_spec.submodule_search_locations)
        return spec


_installed = False


def install():
    global _installed
    if _installed:
        return
    # The reference arm of bench.py puts the genuine paddle first on sys.path and must win.
    import os
    if os.environ.get("PADDLE_B200_NO_ALIAS") == "1":
        return
    sys.meta_path.insert(0, _AliasFinder())
    _installed = True
