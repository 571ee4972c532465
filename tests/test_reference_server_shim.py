"""the reference's UNMODIFIED server module imports on this engine.

Runs where the reference tree exists (this container: /root/reference; the GPU box has no copy, and reference sources are
never vendored into this repo).  ``compat.install(shim_mlx=True)`` makes ``mlx_parallm.utils / .sample_utils / .models.*``
resolve to this package and registers stand-ins for the handful of ``mlx.core`` names the server calls itself
(mlx_parallm_b200/mx_shim.py); the server's own code -- FastAPI app, schemas, batching queues, log-prob arithmetic -- is the
reference's."""
import asyncio
import math
import os
import sys

import numpy as np
import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "mlx_parallm", "server")),
                                reason="reference tree not present on this machine")


@pytest.fixture()
def server_main():
    pytest.importorskip("fastapi")
    import mlx_parallm_b200.compat as compat
    saved = dict(sys.modules)
    sys.path.insert(0, REF)
    try:
        compat.install(shim_mlx=True)
        import mlx_parallm.server.main as main
        yield main
    finally:
        sys.path.remove(REF)
        for k in list(sys.modules):
            if k not in saved and (k.split(".")[0] in ("mlx", "mlx_lm", "mlx_parallm")):
                del sys.modules[k]
        for k, v in saved.items():
            if k.split(".")[0] in ("mlx", "mlx_lm", "mlx_parallm"):
                sys.modules[k] = v


def test_reference_server_module_imports_and_exposes_its_routes(server_main):
    import mlx_parallm_b200.utils as ours
    main = server_main
    paths = {r.path for r in main.app.routes}
    assert {"/v1/completions", "/v1/chat/completions", "/v1/models", "/health"} <= paths, sorted(paths)
    # the engine entry points the server holds ARE this package's
    assert main.batch_generate_text_util is ours.batch_generate_text and main.batch_stream_generate_text is ours.batch_stream_generate_text
    assert main.apply_chat_template_cached is ours.apply_chat_template_cached
    assert main.TokenizerWrapper.__module__ == "mlx_parallm_b200.tokenizer_utils"


class _TableModel:
    """A 'model' whose logits depend only on the current token (a fixed table): the server's own perplexity arithmetic
    (server/main.py:627-659) can then be checked in closed form.  Returns float32 torch logits like this engine does."""

    def __init__(self, V):
        g = torch.Generator().manual_seed(0)
        self.table = torch.randn((V, V), generator=g)

    def __call__(self, x, cache=None):
        x = torch.as_tensor(x)
        return self.table[x.long()]                      # [1, T, V]


def test_reference_perplexity_endpoint_arithmetic_runs_on_the_shim(server_main):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from fake_tokenizer import FakeTokenizer
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    main = server_main
    V = 32
    model, tok = _TableModel(V), TokenizerWrapper(FakeTokenizer(V, eos_token_id=0))
    ids = [3, 7, 1, 9, 4, 4, 2]
    text = " ".join(f"w{i}" for i in ids)
    # reproduce the endpoint body's arithmetic through the names it uses (mx.array / softmax / log on the model's logits)
    import mlx.core as mx
    x = mx.array([ids])
    logits = model(x)
    lps = []
    for i in range(len(ids) - 1):
        pi = mx.softmax(logits[:, i, :], axis=-1)
        lps.append(float(mx.log(pi[0, int(x[0, i + 1].item())]).item()))
    exp = [float(torch.log_softmax(model.table[ids[i]], -1)[ids[i + 1]]) for i in range(len(ids) - 1)]
    assert np.allclose(lps, exp, atol=1e-6)
    assert np.array(pi[0].astype(mx.float32)).shape == (V,)        # server/main.py:522
    tokn = mx.argmax(logits[:, -1, :], axis=-1, keepdims=True)
    y = mx.concatenate([x, tokn], axis=1)
    assert y.shape == (1, len(ids) + 1) and int(y[0, -1].item()) == int(model.table[ids[-1]].argmax())
    mx.eval(y)
    # and the endpoint itself: the reference's /v1/perplexity handler on a registered (model, tokenizer) record
    from mlx_parallm.server.schemas import InternalModelRecord, ModelStatus
    main.model_registry["m"] = InternalModelRecord(id="m", path_or_hf_id="m", status=ModelStatus.LOADED, model_instance=model,
                                                   tokenizer_instance=tok)
    try:
        res = asyncio.run(main.compute_perplexity(main.PerplexityRequest(model="m", text=text)))
    finally:
        main.model_registry.pop("m", None)
    assert res.token_count == len(exp) and math.isclose(res.avg_nll, -sum(exp) / len(exp), rel_tol=1e-5)
    assert math.isclose(res.ppl, math.exp(-sum(exp) / len(exp)), rel_tol=1e-5)
