"""The slice of ``mlx.core`` / ``mlx.nn`` / ``mlx_lm.tokenizer_utils`` that the reference's SERVER touches directly, over torch.

The reference server (mlx_parallm/server/main.py) takes its engine from ``mlx_parallm.utils`` -- replaced by this package
through ``compat.install()`` -- but its log-prob / perplexity endpoints also call MLX themselves on the logits the model
returns (server/main.py:467-653): ``mx.array, mx.softmax, mx.log, mx.argmax, mx.concatenate, mx.eval, mx.float32`` and the
array methods ``.shape, .astype, .item, [...]``; ``nn.Module`` appears in type annotations only (281, 459, 664) and
``mlx_lm.tokenizer_utils.TokenizerWrapper`` is imported by name (32).  ``install()`` registers stand-ins for exactly those names
when the real packages are not importable, so that the unmodified server module imports and its endpoints run on the logits
of this engine (torch CUDA tensors).  It is not an MLX implementation: anything outside this list raises AttributeError.

Arrays are ``torch.Tensor`` subclass instances (so the model accepts them as token ids and indexing keeps the type) with the
three MLX method names the server uses that torch does not have."""
from __future__ import annotations

import sys
import types

import numpy as np
import torch


class array(torch.Tensor):
    """``mx.array``: a torch tensor with ``.astype`` (server/main.py:522, 552, 586)."""

    @staticmethod
    def __new__(cls, data=None, dtype=None):
        if isinstance(data, torch.Tensor):
            t = data
        else:
            t = torch.as_tensor(np.asarray(data))
        if dtype is not None:
            t = t.to(dtype)
        return t.as_subclass(cls)

    def astype(self, dtype):
        return self.to(dtype)

    def __array__(self, dtype=None, copy=None):
        a = self.detach().as_subclass(torch.Tensor).cpu().numpy()
        return a.astype(dtype) if dtype is not None else a


def _wrap(t):
    return t.as_subclass(array) if isinstance(t, torch.Tensor) and not isinstance(t, array) else t


def softmax(x, axis=-1):
    """mx.softmax: computed in float32 like MLX's precise path for 16-bit inputs is NOT assumed -- the engine's logits are
    float32 already."""
    return _wrap(torch.softmax(torch.as_tensor(x).float(), dim=axis))


def log(x):
    return _wrap(torch.log(torch.as_tensor(x)))


def argmax(x, axis=None, keepdims=False):
    x = torch.as_tensor(x)
    if axis is None:
        return _wrap(torch.argmax(x))
    return _wrap(torch.argmax(x, dim=axis, keepdim=keepdims))       # lowest index among ties, like mx.argmax on CUDA


def concatenate(arrays, axis=0):
    ts = [torch.as_tensor(a) for a in arrays]
    dev = next((t.device for t in ts if t.is_cuda), ts[0].device)
    return _wrap(torch.cat([t.to(dev) for t in ts], dim=axis))


def eval(*arrays):
    """mx.eval forces lazy graphs; torch is eager -- wait for the stream so that timing code around it keeps its meaning."""
    if torch.cuda.is_available():
        torch.cuda.current_stream().synchronize()


def _core_module() -> types.ModuleType:
    m = types.ModuleType("mlx.core")
    m.__doc__ = __doc__
    m.array, m.softmax, m.log, m.argmax, m.concatenate, m.eval = array, softmax, log, argmax, concatenate, eval
    m.float32, m.float16, m.bfloat16, m.int32, m.int64 = torch.float32, torch.float16, torch.bfloat16, torch.int32, torch.int64
    return m


def _nn_module() -> types.ModuleType:
    m = types.ModuleType("mlx.nn")

    class Module:                                   # annotation target only (server/main.py:281, 459, 664)
        pass

    m.Module = Module
    return m


def install(force: bool = False) -> bool:
    """Register the stand-ins unless the real ``mlx`` is importable (then nothing is touched).  Returns True if installed."""
    if not force:
        try:
            import mlx.core  # noqa: F401
            return False
        except Exception:
            pass
    from . import tokenizer_utils
    mlx = types.ModuleType("mlx")
    mlx.__path__ = []
    mlx.core, mlx.nn = _core_module(), _nn_module()
    mlx_lm = types.ModuleType("mlx_lm")
    mlx_lm.__path__ = []
    mlx_lm.tokenizer_utils = tokenizer_utils
    sys.modules.update({"mlx": mlx, "mlx.core": mlx.core, "mlx.nn": mlx.nn, "mlx_lm": mlx_lm,
                        "mlx_lm.tokenizer_utils": tokenizer_utils})
    return True
