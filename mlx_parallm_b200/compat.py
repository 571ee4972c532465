"""Drop-in aliasing of the reference's module names.

    import mlx_parallm_b200.compat as compat
    compat.install()                       # before anything imports mlx_parallm.utils & co
    from mlx_parallm.utils import load, generate_step, batch_generate        # -> this engine

After ``install()`` the names the reference's callers import (server/main.py:22-33, demo.py:5) resolve to this package:
``mlx_parallm.utils``, ``mlx_parallm.sample_utils``, ``mlx_parallm.models``, ``mlx_parallm.models.base / llama / qwen3``.
If a real ``mlx_parallm`` package is importable its other submodules (server, cli, rl_training) stay reachable -- only the
engine modules are replaced."""
from __future__ import annotations

import importlib
import sys
import types

ENGINE_MODULES = ("utils", "sample_utils", "tokenizer_utils", "models", "models.base", "models.llama", "models.qwen3")


def install(force: bool = True, shim_mlx: bool = False) -> types.ModuleType:
    """``shim_mlx``: also register the stand-ins of mx_shim.py for the ``mlx.core`` / ``mlx.nn`` / ``mlx_lm.tokenizer_utils``
    names the reference SERVER uses itself (only when the real packages are not importable), so that
    ``import mlx_parallm.server.main`` works from a reference checkout on sys.path."""
    if shim_mlx:
        from . import mx_shim
        mx_shim.install()
    pkg = sys.modules.get("mlx_parallm")
    if pkg is None:
        try:
            pkg = importlib.import_module("mlx_parallm")
        except Exception:                       # not installed, or its own imports (mlx) fail: create the namespace
            pkg = types.ModuleType("mlx_parallm")
            pkg.__path__ = []                   # a package, so "import mlx_parallm.x" consults sys.modules
            sys.modules["mlx_parallm"] = pkg
    for name in ENGINE_MODULES:
        full = f"mlx_parallm.{name}"
        if full in sys.modules and not force:
            continue
        mod = importlib.import_module(f"mlx_parallm_b200.{name}")
        sys.modules[full] = mod
        parent, _, leaf = full.rpartition(".")
        setattr(sys.modules[parent], leaf, mod)
    return pkg
