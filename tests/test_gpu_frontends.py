"""The reference's text front-ends on the CUDA engine: batch_generate, batch_generate_text (LCP split,
shared-prefix prefill), batch_stream_generate_text, generate / stream_generate, the KV pool and two threads on one model.

Expected transcripts come from the CPU oracle fed the SAME padded ids (left pads are attended and take RoPE positions,
SURVEY B1), with the reference's bookkeeping (utils.py:520-531, 1028-1068, 1296-1340) restated on the oracle's tokens."""
import asyncio
import threading

import numpy as np
import pytest
import torch

from fake_tokenizer import FakeTokenizer
from helpers import make_pair
from oracle.model import generate_step as oracle_generate_step

pytestmark = pytest.mark.gpu

NAME, DTYPE = "small-llama", "float16"


@pytest.fixture(scope="module")
def pair():
    cfg, om, gm = make_pair(NAME, dtype=DTYPE)
    return cfg, om, gm


def oracle_tokens(om, ids, steps, cache=None):
    return np.stack([y[:, 0] for _, (y, _) in zip(range(steps), oracle_generate_step(np.asarray(ids), om, cache=cache))], 1)


def words(ids):
    return " ".join(f"w{int(i)}" for i in ids)


def left_pad(rows, pad):
    n = max(len(r) for r in rows)
    return np.array([[pad] * (n - len(r)) + list(r) for r in rows], dtype=np.int64)


# Fixture prompts are chosen so that every greedy decision of the oracle has a top-2 margin >= 3e-3 of max|logits| (the
# engine sits ~4e-4 away, and its fp32 accumulate epilogue sums stream-K partials in arrival order, so a decision closer
# than that may flip between two runs): re-check with the oracle when changing them.
def pick_eos(tokens, step):
    """An id that row 0 emits at ``step`` and no row emits earlier: used as EOS it ends row 0 mid-batch."""
    cand = int(tokens[0, step])
    assert not (tokens[:, :step] == cand).any(), "fixture prompts changed: choose another EOS step"
    return cand


# ------------------------------------------------------------------------------------------ batch_generate (a13)
def test_batch_generate_left_pads_and_matches_the_oracle(pair):
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    from mlx_parallm_b200.utils import batch_generate
    cfg, om, gm = pair
    rows = [[38, 39, 40, 41, 42, 43, 44], [148, 149, 150], [258, 259, 260, 261, 262]]
    prompts = [words(r) for r in rows]
    # no chat template: the prompt batch is the rows left-padded with eos (pad unset -> eos, utils.py:513-515)
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=0))
    got = batch_generate(gm, tok, prompts, max_tokens=12, format_prompts=False)
    exp_t = oracle_tokens(om, left_pad(rows, 0), 12)
    exp = [words(t).split("<eos>")[0] for t in exp_t]
    assert [g.strip() for g in got] == [e.strip() for e in exp]
    assert tok._tokenizer.padding_side == "left" and tok._tokenizer.pad_token == "<eos>"
    # with the chat template: 'w1 <prompt> w2' per row (format_prompts=True is the default)
    got = batch_generate(gm, tok, prompts, max_tokens=6)
    exp_t = oracle_tokens(om, left_pad([[1] + r + [2] for r in rows], 0), 6)
    assert [g.strip() for g in got] == [words(t).strip() for t in exp_t]


def test_batch_generate_runs_max_tokens_steps_and_cuts_at_the_eos_string(pair):
    """No early stop on EOS: every row runs max_tokens steps, the text is cut at the eos string afterwards."""
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    from mlx_parallm_b200.utils import batch_generate
    cfg, om, gm = pair
    rows = [[41, 42, 43, 44], [51, 52, 53, 54]]
    t0 = oracle_tokens(om, np.array(rows), 10)
    eos = pick_eos(t0, 4)
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=eos))
    got = batch_generate(gm, tok, [words(r) for r in rows], max_tokens=10, format_prompts=False)
    assert got[0].strip() == words(t0[0, :4])                      # cut where row 0 emitted the eos id
    other = list(t0[1])
    exp1 = words(other[:other.index(eos)]) if eos in other else words(other)
    assert got[1].strip() == exp1


# ------------------------------------------------------------------------------------------ batch_generate_text (a14)
def _bgt(gm, tok, prompts, **kw):
    from mlx_parallm_b200.utils import batch_generate_text
    return asyncio.run(batch_generate_text(gm, tok, prompts, **kw))


@pytest.mark.parametrize("disable_prefix_cache", [False, True])
def test_batch_generate_text_shared_prefix_equal_lengths(pair, disable_prefix_cache):
    """Equal-length prompts with a common prefix: LCP prefill + suffix batch is the same computation as one prefill of the
    full prompts (no pads anywhere), with or without the prefix split."""
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    cfg, om, gm = pair
    prefix = [100, 101, 102, 103, 104, 105]
    rows = [prefix + [201, 202, 203], prefix + [301, 302, 303], prefix + [401, 402, 403]]
    t0 = oracle_tokens(om, np.array(rows), 10)
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=0))
    out = _bgt(gm, tok, [words(r) for r in rows], max_tokens=10, temp=0.0, disable_prefix_cache=disable_prefix_cache)
    for i, (text, n_prompt, n_comp) in enumerate(out):
        row = list(t0[i])
        exp = row[:row.index(0)] if 0 in row else row
        assert text == words(exp) and n_prompt == len(rows[i]) and n_comp == len(exp)


def test_batch_generate_text_ragged_suffixes_after_the_shared_prefix(pair):
    """Ragged prompts: the shared prefix is prefilled once, the suffixes are LEFT-padded with eos behind it (utils.py:1281-
    1290) -- pads are real tokens at positions lcp.. -- and the oracle fed prefix + padded suffix gives the same tokens."""
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    cfg, om, gm = pair
    prefix = [100, 101, 102, 103]
    rows = [prefix + [201, 202, 203, 204], prefix + [301], prefix + [401, 402]]
    suffix = left_pad([r[len(prefix):] for r in rows], 0)
    full = np.concatenate([np.tile(np.array(prefix), (3, 1)), suffix], 1)
    t0 = oracle_tokens(om, full, 8)
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=0))
    out = _bgt(gm, tok, [words(r) for r in rows], max_tokens=8, temp=0.0)
    for i, (text, n_prompt, n_comp) in enumerate(out):
        row = list(t0[i])
        exp = row[:row.index(0)] if 0 in row else row
        assert text == words(exp) and n_prompt == len(rows[i]) and n_comp == len(exp)


def test_batch_generate_text_single_prompt_eos_and_max_tokens(pair):
    """One prompt: lcp = the whole prompt, the suffix batch is a single pad token (utils.py:1225, 1288-1290); EOS ends the
    row, max_tokens caps it."""
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    cfg, om, gm = pair
    row = [64, 65, 66, 67, 68]
    t_probe = oracle_tokens(om, np.array([row + [0]]), 12)          # prompt + the pad token the front-end appends
    eos = pick_eos(t_probe, 4)
    t0 = oracle_tokens(om, np.array([row + [eos]]), 12)             # the pad IS eos: recompute with it in place
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=eos))
    (text, n_prompt, n_comp), = _bgt(gm, tok, [words(row)], max_tokens=12, temp=0.0)
    r = list(t0[0])
    exp = r[:r.index(eos)] if eos in r else r
    assert text == words(exp) and n_prompt == len(row) and n_comp == len(exp)
    (text, _, n_comp), = _bgt(gm, tok, [words(row)], max_tokens=3, temp=0.0)
    assert n_comp == min(3, len(exp)) and text == words(exp[:3])
    assert _bgt(gm, tok, [], max_tokens=3) == []


# ------------------------------------------------------------------------------------------ streaming
def test_batch_stream_generate_text_deltas_and_finish_reasons(pair):
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    from mlx_parallm_b200.utils import batch_stream_generate_text
    cfg, om, gm = pair
    rows = np.array([[80, 81, 82, 83], [190, 191, 192, 193], [300, 301, 302, 303]])
    t0 = oracle_tokens(om, rows, 9)
    eos = pick_eos(t0, 3)
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=eos))
    text = ["", "", ""]
    reason = [None, None, None]
    steps_seen = 0
    for deltas in batch_stream_generate_text(gm, tok, rows, max_tokens=9):
        steps_seen += 1
        assert len(deltas) == 3
        for i, (d, r) in enumerate(deltas):
            if reason[i] is not None:
                assert (d, r) == (None, None)                      # a finished row stays silent
                continue
            if d:
                text[i] += d
            reason[i] = r
    for i in range(3):
        row = list(t0[i])
        if eos in row:
            assert reason[i] == "stop" and text[i].strip() == words(row[:row.index(eos)])
        else:
            assert reason[i] == "length" and text[i].strip() == words(row)
    assert reason[0] == "stop" and steps_seen <= 9


def test_generate_and_stream_generate_single_prompt(pair):
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    from mlx_parallm_b200.utils import generate, stream_generate
    cfg, om, gm = pair
    row = [15, 25, 35, 45, 55, 65]
    t0 = oracle_tokens(om, np.array([row]), 10)
    eos = pick_eos(t0, 5)
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=eos))
    assert generate(gm, tok, words(row), max_tokens=10).strip() == words(t0[0, :5])
    assert "".join(stream_generate(gm, tok, words(row), max_tokens=10)).strip() == words(t0[0, :5])
    assert generate(gm, tok, words(row), max_tokens=4).strip() == words(t0[0, :4])
    with pytest.raises(NotImplementedError):
        generate(gm, tok, words(row), max_tokens=2, repetition_penalty=1.1)


# ------------------------------------------------------------------------------------------ pool (a12), threads
def test_kv_pool_reuses_buffers_and_the_decode_graph(pair):
    """Same thread, same model, same batch size -> the same cache objects, buffers and captured decode graph come back reset;
    another batch size gets its own entry; results do not depend on what ran before."""
    from mlx_parallm_b200 import utils
    cfg, om, gm = pair
    kvh = [l.self_attn.n_kv_heads for l in gm.layers]
    a = np.array([[124, 125, 126, 127, 128], [174, 175, 176, 177, 178], [224, 225, 226, 227, 228]])
    b = np.array([[124, 125, 126], [184, 185, 186]])

    def run(ids, n=6):
        return np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(n), utils.generate_step(torch.from_numpy(ids), gm))], 1)

    ta = run(a)
    ca = utils._kv_pool.get(gm.head_dim, kvh, 3, device=gm.device, owner=id(gm))
    ptr, graphs = ca[0].keys.data_ptr(), dict(ca[0].__dict__.get("_decode_steps", {}))
    assert graphs and ca[0].offsets == [0, 0, 0]                  # get() hands the entry back reset
    tb = run(b)
    cb = utils._kv_pool.get(gm.head_dim, kvh, 2, device=gm.device, owner=id(gm))
    assert cb is not ca and cb[0].batch_size == 2
    ta2 = run(a)
    ca2 = utils._kv_pool.get(gm.head_dim, kvh, 3, device=gm.device, owner=id(gm))
    assert ca2 is ca and ca2[0].keys.data_ptr() == ptr            # same buffers
    st_before, st_after = list(graphs.values())[0], list(ca2[0].__dict__["_decode_steps"].values())[0]
    assert st_after is st_before and st_after.graph is not None   # same captured graph, replayed
    assert np.array_equal(ta, ta2)
    assert np.array_equal(ta, oracle_tokens(om, a, 6)) and np.array_equal(tb, oracle_tokens(om, b, 6))


def test_two_threads_generate_on_one_model(pair):
    """SURVEY 8b 'Threading': the reference server runs generate_step from an executor thread and from the event loop at once
    (utils.py:1345 vs server/main.py:1358).  Same batch size on purpose: both threads use the same scratch arena."""
    from mlx_parallm_b200 import utils
    cfg, om, gm = pair
    prompts = [np.array([[43, 44, 45, 46], [53, 54, 55, 56]]), np.array([[93, 94, 95, 96], [103, 104, 105, 106]])]
    exp = [oracle_tokens(om, p, 20) for p in prompts]
    got, errs = [None, None], []
    start = threading.Barrier(2)

    def worker(i):
        try:
            torch.cuda.set_device(0)
            start.wait()
            toks = []
            for _, (y, _) in zip(range(20), utils.generate_step(torch.from_numpy(prompts[i]), gm)):
                toks.append(y[:, 0].cpu().numpy())
            got[i] = np.stack(toks, 1)
        except Exception as e:                                     # surfaces in the main thread
            errs.append(e)

    ths = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(120)
    assert not errs, errs
    for i in range(2):
        assert np.array_equal(got[i], exp[i]), i


def test_concurrent_batch_generate_text_calls(pair):
    """Two batch_generate_text coroutines at once: both run _synchronous_generation in executor threads on ONE model."""
    from mlx_parallm_b200.tokenizer_utils import TokenizerWrapper
    from mlx_parallm_b200.utils import batch_generate_text
    cfg, om, gm = pair
    tok = TokenizerWrapper(FakeTokenizer(cfg["vocab_size"], eos_token_id=0))
    sets = [[[7, 8, 9, 10], [17, 18, 19, 20]], [[27, 28, 29, 30], [37, 38, 39, 40]]]

    async def both():
        return await asyncio.gather(*[batch_generate_text(gm, tok, [words(r) for r in rows], max_tokens=8, temp=0.0)
                                      for rows in sets])

    outs = asyncio.run(both())
    for rows, out in zip(sets, outs):
        t0 = oracle_tokens(om, np.array(rows), 8)
        for i, (text, n_prompt, n_comp) in enumerate(out):
            r = list(t0[i])
            exp = r[:r.index(0)] if 0 in r else r
            assert text == words(exp) and n_prompt == 4 and n_comp == len(exp)
