"""Shared test helpers: build the same synthetic model for the oracle (numpy) and the CUDA path."""
import numpy as np
import torch

from mlx_parallm_b200 import synth
from oracle.model import OracleConfig, OracleModel


def make_pair(name, dtype="float16", seed=1234, numerics="reference", device="cuda", **overrides):
    """-> (cfg, oracle model, CUDA model or None if no device requested)."""
    cfg = synth.get_config(name, **overrides)
    w = synth.random_weights(cfg, seed=seed, dtype=dtype, device="cpu")
    ocfg = OracleConfig.from_hf(cfg, dtype)
    om = OracleModel(ocfg, synth.to_numpy_weights(w), numerics=numerics)
    gm = None
    if device is not None:
        from mlx_parallm_b200.utils import _get_classes
        Model, Args = _get_classes(cfg)
        gm = Model(Args.from_dict(cfg)).load_weights(w, device=device)
    return cfg, om, gm


def rel_err(a, b):
    """max |a-b| / max |b| -- the 'relative' of the north star's 1e-2 logits tolerance."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def prompts_for(cfg, B, L, seed=7):
    return np.random.default_rng(seed).integers(0, cfg["vocab_size"], (B, L)).astype(np.int64)


def greedy_report(got, exp, exp_logits, tol_rel):
    """got / exp: [B, S] token ids; exp_logits: list of S arrays [B, V] (the oracle's logits per step).
    Returns (identical fraction, near-tie divergences, real divergences).  A sequence is followed up to its first
    divergence (after it the two contexts differ and nothing can be compared); the divergence is a near-tie when the oracle's
    logit of the engine's token is within tol_rel * max|logits| of the oracle's maximum."""
    B, S = exp.shape
    identical, near, real = 0, 0, []
    for b in range(B):
        d = np.nonzero(got[b] != exp[b])[0]
        if d.size == 0:
            identical += 1
            continue
        s = int(d[0])
        lg = exp_logits[s][b]
        margin = float(lg.max() - lg[got[b, s]])
        if margin <= tol_rel * float(np.abs(exp_logits[s]).max()):
            near += 1
        else:
            real.append((b, s, margin))
    return identical / B, near, real
