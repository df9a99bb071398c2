"""paddle.hub (reference: python/paddle/hapi/hub.py): entry points published by a repository's ``hubconf.py``.

``source="local"`` reads ``<repo_dir>/hubconf.py``.  ``github`` / ``gitee`` repositories are used from the cache
directory (``~/.cache/paddle/hub/<owner>_<repo>_<branch>``) when a copy is already there; fetching one needs a network,
which this environment does not have, and is reported as such."""
from __future__ import annotations

import importlib.util
import os
import sys

__all__ = ["list", "help", "load"]

VAR_DEPENDENCY = "dependencies"
MODULE_HUBCONF = "hubconf.py"
HUB_DIR = os.path.expanduser(os.path.join("~", ".cache", "paddle", "hub"))
_builtin_list = list


def _import_hubconf(repo_dir):
    path = os.path.join(repo_dir, MODULE_HUBCONF)
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} not found: a hub repository needs a {MODULE_HUBCONF} at its root")
    sys.path.insert(0, repo_dir)
    try:
        spec = importlib.util.spec_from_file_location(MODULE_HUBCONF[:-3], path)
        module = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(module)
    finally:
        sys.path.remove(repo_dir)
    return module


def _check_dependencies(module):
    deps = getattr(module, VAR_DEPENDENCY, None)
    if deps is not None:
        missing = [pkg for pkg in deps if importlib.util.find_spec(pkg) is None]
        if missing:
            raise RuntimeError("Missing dependencies: {}".format(", ".join(missing)))


def _resolve(repo_dir, source, force_reload):
    if source not in ("github", "gitee", "local"):
        raise ValueError(f'Unknown source: "{source}". Allowed values: "github" | "gitee" | "local".')
    if source == "local":
        return repo_dir
    branch = "main" if source == "github" else "master"
    repo = repo_dir
    if ":" in repo:
        repo, branch = repo.split(":")
    owner, name = repo.split("/")
    cached = os.path.join(HUB_DIR, "_".join([owner, name, branch.replace("/", "_")]))
    if os.path.isdir(cached) and not force_reload:
        return cached
    raise RuntimeError(f"{source} repository {repo_dir} is not in the cache ({cached}) and cannot be downloaded: this environment "
                       f"has no network access; clone it and use source='local'")


def _entry(module, name):
    if not isinstance(name, str):
        raise ValueError("Invalid input: model should be a str of function name")
    fn = getattr(module, name, None)
    if fn is None or not callable(fn):
        raise RuntimeError(f"Cannot find callable {name} in hubconf")
    return fn


def list(repo_dir, source="github", force_reload=False):  # noqa: A001
    """Names of the public callables of the repository's hubconf."""
    module = _import_hubconf(_resolve(repo_dir, source, force_reload))
    return [f for f in dir(module) if callable(getattr(module, f)) and not f.startswith("_")]


def help(repo_dir, model, source="github", force_reload=False):  # noqa: A001
    """Docstring of entry point ``model``."""
    module = _import_hubconf(_resolve(repo_dir, source, force_reload))
    return _entry(module, model).__doc__


def load(repo_dir, model, source="github", force_reload=False, **kwargs):
    """Call entry point ``model`` with ``kwargs`` (after checking the hubconf's ``dependencies`` list)."""
    module = _import_hubconf(_resolve(repo_dir, source, force_reload))
    _check_dependencies(module)
    return _entry(module, model)(**kwargs)
