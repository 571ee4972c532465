"""TEST INFRASTRUCTURE -- the hook that pins this oracle against the real reference the day MLX is importable.

The reference's arithmetic lives in ``mlx==0.25.2`` / ``mlx-lm==0.24.1`` (pyproject.toml:38-39), which are installed on
neither this container nor the GPU box (no wheel, no network), so parity is "unpinned" today.  This
module is what closes that gap without a code change: when ``import mlx.core`` works AND the reference tree is on disk,
``available()`` turns true and ``tests/test_mlx_adapter.py`` stops skipping.  It then

  * runs the reference's OWN ``load_model`` (mlx_parallm/utils.py:630-708) and ``generate_step`` (utils.py:315-427) on
    ``mx.cpu`` over a synthetic checkpoint written by ``mlx_parallm_b200.synth.save_checkpoint`` (same safetensors, same
    config.json the engine loads);
  * exposes ``mx.dequantize`` on raw (codes, scales, biases) so the CUDA kernel's rounding recipe (``B200P_DQ_*``,
    include/b200p.h) can be locked bit for bit;
  * reports ``oracle_kind`` = "mlx-cpu-<version>" (else "restatement").

Nothing here is imported by the product; only tests/ may use it (tests/test_abi.py enforces that for all of oracle/)."""
from __future__ import annotations

import os
import sys
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np

REFERENCE_ROOT = os.environ.get("MLX_PARALLM_REFERENCE", "/root/reference")


def available() -> bool:
    """True when real MLX is importable and the reference package is on disk."""
    if not os.path.isdir(os.path.join(REFERENCE_ROOT, "mlx_parallm")):
        return False
    try:
        import mlx.core as mx  # noqa: F401
    except Exception:
        return False
    # the stand-ins of mlx_parallm_b200/mx_shim.py are not MLX
    return not getattr(sys.modules.get("mlx.core"), "__doc__", "").startswith("The slice of")


def oracle_kind() -> str:
    if not available():
        return "restatement"
    import mlx.core as mx
    return f"mlx-cpu-{getattr(mx, '__version__', 'unknown')}"


def _reference_utils():
    """The reference's own utils module, imported from its tree (never the compat alias of this engine)."""
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    for k in [k for k in sys.modules if k == "mlx_parallm" or k.startswith("mlx_parallm.")]:
        if "mlx_parallm_b200" in getattr(sys.modules[k], "__file__", "") or "":
            del sys.modules[k]                      # drop compat aliases so the real package is imported
    import importlib
    return importlib.import_module("mlx_parallm.utils")


def dequantize(codes_u32: np.ndarray, scales: np.ndarray, biases: np.ndarray, group_size: int = 64, bits: int = 4) -> np.ndarray:
    """``mx.dequantize`` on the CPU device; result in the scales' dtype, returned as numpy (bf16 comes back as float32
    holding bf16 values)."""
    import mlx.core as mx
    mx.set_default_device(mx.cpu)
    dt = {"float16": mx.float16, "float32": mx.float32}.get(str(scales.dtype), mx.bfloat16)
    w = mx.array(np.ascontiguousarray(codes_u32).view(np.uint32))
    s = mx.array(np.asarray(scales, np.float32)).astype(dt)
    b = mx.array(np.asarray(biases, np.float32)).astype(dt)
    out = mx.dequantize(w, s, b, group_size=group_size, bits=bits)
    mx.eval(out)
    return np.array(out.astype(mx.float32))


def quantize(w: np.ndarray, group_size: int = 64, bits: int = 4) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """``mx.quantize`` -> (codes uint32, scales, biases as float32 numpy holding the storage-type values)."""
    import mlx.core as mx
    mx.set_default_device(mx.cpu)
    dt = {"float16": mx.float16, "float32": mx.float32}.get(str(w.dtype), mx.bfloat16)
    q, s, b = mx.quantize(mx.array(np.asarray(w, np.float32)).astype(dt), group_size=group_size, bits=bits)
    mx.eval(q, s, b)
    return np.array(q), np.array(s.astype(mx.float32)), np.array(b.astype(mx.float32))


class ReferenceRun:
    """The reference's model on ``mx.cpu`` loaded from a checkpoint directory (config.json + model.safetensors)."""

    def __init__(self, checkpoint_dir: str):
        import mlx.core as mx
        from pathlib import Path
        mx.set_default_device(mx.cpu)
        self.mx = mx
        self.utils = _reference_utils()
        self.model = self.utils.load_model(Path(checkpoint_dir))

    def logits(self, ids: np.ndarray) -> np.ndarray:
        """All-position logits of a fresh forward (no cache): [B, L, V] float32."""
        mx = self.mx
        out = self.model(mx.array(np.asarray(ids)))
        mx.eval(out)
        return np.array(out.astype(mx.float32))

    def greedy(self, prompts: np.ndarray, steps: int) -> Iterator[Tuple[np.ndarray, Optional[np.ndarray]]]:
        """The reference's ``generate_step`` (temp 0): yields token ids [B] per step."""
        mx = self.mx
        gen = self.utils.generate_step(mx.array(np.asarray(prompts)), self.model, temp=0.0)
        for _, (tok, _) in zip(range(steps), gen):
            mx.eval(tok)
            yield np.array(tok).reshape(-1), None


def compare_all(checkpoint_dir: str, prompts: np.ndarray, steps: int = 32) -> Dict[str, object]:
    """Everything SURVEY 8c asks of the adapter in one call (used by tests/test_mlx_adapter.py): reference logits and
    greedy transcript for ``prompts`` from the real MLX path."""
    run = ReferenceRun(checkpoint_dir)
    toks: List[np.ndarray] = [t for t, _ in run.greedy(prompts, steps)]
    return {"oracle_kind": oracle_kind(), "logits": run.logits(prompts), "tokens": np.stack(toks, 1)}
