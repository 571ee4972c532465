"""Generation engine with the reference's public surface (mlx_parallm/utils.py).

    load / load_model / load_config           utils.py:630-747
    generate_step                             utils.py:315-427
    generate / stream_generate                utils.py:429-471, 545-616
    batch_generate                            utils.py:473-543
    batch_stream_generate_text                utils.py:983-1075
    batch_generate_text (async)               utils.py:1087-1346
    _KVPool                                   utils.py:199-226

B200 design: prefill runs the layer kernels eagerly in chunks; the decode step (embed -> L blocks ->
head -> argmax -> lengths+=1) is captured ONCE per (batch, cache buffers) into a CUDA graph and
replayed per token with the sampled ids fed back on the device, so a step costs one graph launch and no
host<->device sync; callers that read tokens (``.item()``/``.tolist()``) pay their own D2H.
"""
from __future__ import annotations

import asyncio
import copy
import gc
import glob
import importlib
import json
import logging
import threading
import time
from collections import OrderedDict
from pathlib import Path
from typing import Any, Callable, Dict, Generator, List, Optional, Tuple, Union

import numpy as np
import torch

from . import ops, sample_utils
from .models._transformer import DecoderModel, advance_cache, bind_cache_group, prepare_cache
from .models.base import BatchedKVCache, PagedKVCache
from .sample_utils import top_p_sampling
from .tokenizer_utils import TokenizerWrapper, load_tokenizer

MODEL_REMAPPING = {"mistral": "llama"}        # utils.py:33-36


class ModelNotFoundError(Exception):
    def __init__(self, message):
        self.message = message
        super().__init__(self.message)


def _get_classes(config: dict):
    """utils.py:47-67."""
    model_type = MODEL_REMAPPING.get(config["model_type"], config["model_type"])
    try:
        arch = importlib.import_module(f"mlx_parallm_b200.models.{model_type}")
    except ImportError:
        msg = f"Model type {model_type} not supported."
        logging.error(msg)
        raise ValueError(msg)
    return arch.Model, arch.ModelArgs


def get_model_path(path_or_hf_repo: str, revision: Optional[str] = None) -> Path:
    """utils.py:70-104 without the hub download (no network): the path must exist locally."""
    p = Path(path_or_hf_repo)
    if not p.exists():
        raise ModelNotFoundError(f"Model not found for path or HF repo: {path_or_hf_repo} (offline build: local paths only)")
    return p


# --------------------------------------------------------------------------------- tokenisation / chat-template caches
class _LRUCache:
    """utils.py:137-156: the small LRU behind encode_cached / apply_chat_template_cached (pure host code, no MLX)."""

    def __init__(self, max_size: int = 2048):
        self.max_size = max_size
        self._store: "OrderedDict[str, Any]" = OrderedDict()
        self._mu = threading.Lock()            # the server calls these from the event loop and from executor threads

    def get(self, key: str):
        with self._mu:
            if key in self._store:
                self._store.move_to_end(key)
                return self._store[key]
            return None

    def set(self, key: str, value: Any):
        with self._mu:
            self._store[key] = value
            self._store.move_to_end(key)
            if len(self._store) > self.max_size:
                self._store.popitem(last=False)


_encode_lru = _LRUCache(max_size=4096)
_chat_template_lru = _LRUCache(max_size=2048)


def encode_cached(tokenizer: TokenizerWrapper, text: str) -> List[int]:
    """utils.py:158-172: LRU-cached ``tokenizer.encode`` for single-string prompts, keyed by tokenizer identity + text."""
    key = f"tok:{id(tokenizer._tokenizer)}:{text}"
    val = _encode_lru.get(key)
    if val is not None:
        return val
    enc = tokenizer.encode(text)
    _encode_lru.set(key, enc)
    return enc


def _messages_cache_key(messages: List[Dict[str, Any]]) -> str:
    """utils.py:174-178: canonical JSON of (role, content) only."""
    minimal = [{"role": m.get("role"), "content": m.get("content")} for m in messages]
    return json.dumps(minimal, sort_keys=True, separators=(",", ":"))


def apply_chat_template_cached(tokenizer: TokenizerWrapper, messages: List[Dict[str, Any]], *,
                               add_generation_prompt: bool = True) -> str:
    """utils.py:180-197 (imported by server/main.py:33): LRU-cached chat templating."""
    key = f"cht:{id(tokenizer._tokenizer)}:{'1' if add_generation_prompt else '0'}:{_messages_cache_key(messages)}"
    val = _chat_template_lru.get(key)
    if val is not None:
        return val
    text = tokenizer.apply_chat_template(messages, tokenize=False, add_generation_prompt=add_generation_prompt)
    _chat_template_lru.set(key, text)
    return text


# --------------------------------------------------------------------------------- KV pool
class _KVPool:
    """utils.py:199-223.  Keyed additionally by model identity and thread so two generations never share
    cache objects (the reference's pool is a documented race)."""

    def __init__(self):
        self._pool: Dict[tuple, List[BatchedKVCache]] = {}
        self._mu = threading.Lock()

    def get(self, head_dim: int, kv_heads: List[int], batch_size: int, *, step: Optional[int] = None,
            paged: bool = True, device=None, owner: int = 0) -> List[BatchedKVCache]:
        """``step`` sizes the FIRST allocation of a new entry and the growth quantum of a reused one; a reused entry keeps
        its (possibly larger) buffers, so the decode graph captured over them stays valid across calls."""
        key = (head_dim, tuple(kv_heads), batch_size, owner, threading.get_ident())
        with self._mu:
            caches = self._pool.get(key)
            if caches is None:
                klass = PagedKVCache if paged else BatchedKVCache
                caches = [klass(head_dim, n, batch_size) for n in kv_heads]
                bind_cache_group(caches, device, step)
                if len(self._pool) > 16:
                    self._pool.clear()
                self._pool[key] = caches
                return caches
        for c in caches:
            c.reset(batch_size)
            if step is not None:
                c.step = step
        return caches


def _presized_cache(model: DecoderModel, B: int, total_tokens: int) -> List[BatchedKVCache]:
    """Pool entry whose buffers take ``total_tokens`` per row in ONE allocation (rounded to 256 like utils.py:1234), so a
    generation of known length never re-allocates or copies its KV cache and its decode graph is captured once."""
    step = max(256, ((int(total_tokens) + 255) // 256) * 256)
    kv_heads = [l.self_attn.n_kv_heads for l in model.layers]
    return _kv_pool.get(model.head_dim, kv_heads, B, step=step, paged=True, device=model.device, owner=id(model))


def _presize(kwargs: dict, model: DecoderModel, prompt_tokens, max_tokens: int) -> dict:
    """generate_step kwargs with a cache that already holds prompt + max_tokens positions (unless the caller brought one)."""
    if kwargs.get("cache") is None and isinstance(model, DecoderModel):
        B, L = prompt_tokens.shape
        kwargs = dict(kwargs, cache=_presized_cache(model, B, L + max_tokens))
    return kwargs


_kv_pool = _KVPool()


# --------------------------------------------------------------------------------- decode graph
class _DecodeStep:
    """One decode step as a replayable CUDA graph over static buffers."""

    def __init__(self, model: DecoderModel, cache: List[BatchedKVCache], B: int, greedy: bool, use_graph: bool = True):
        self.model, self.cache, self.B, self.greedy = model, cache, B, greedy
        dev = model.device
        self.tok = torch.zeros((B,), dtype=torch.int32, device=dev)
        self.probs = torch.zeros((B,), dtype=torch.float32, device=dev)
        head = model.lm_head if model.lm_head is not None else model.model.embed_tokens
        self.logits = torch.empty((B, head.out_features), dtype=torch.float32, device=dev)
        self.arena = model._arena(B, B)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.use_graph = use_graph and model.graph_safe(B)
        self._sig = None

    def _signature(self):
        return tuple((c.keys.data_ptr(), c.values.data_ptr(), c._lengths.data_ptr()) for c in self.cache)

    def _body(self):
        m = self.model
        m._layers_forward(self.tok, self.B, 1, self.cache, self.arena)
        if self.arena.get("x_final"):          # row-sharded TP exchange already produced final_norm(h) in x
            m._head(None, self.arena, logits_out=self.logits, x_rows=self.arena["x"])
        else:
            m._head(self.arena["h"], self.arena, logits_out=self.logits)
        seen = set()
        for c in self.cache:
            if c._lengths.data_ptr() not in seen:
                seen.add(c._lengths.data_ptr())
                ops.advance_lengths(c._lengths, 1)
        if self.greedy:
            m.greedy_tokens(self.logits, self.tok, self.probs)

    def run(self):
        """Consume self.tok (ids of the previous step), leave logits (+ greedy tokens) for this one.  Runs under the
        model's forward lock: the scratch arena and the GEMM workspace belong to one step at a time, and a capture must
        not see another thread's launches (capture_error_mode thread_local: other threads may allocate meanwhile)."""
        with self.model._lock:
            moved = prepare_cache(self.cache, self.B, 1, self.model.dtype, self.model.device)
            arena = self.model._arena(self.B, self.B)     # same object unless the model dropped / re-keyed its arenas
            stale = arena is not self.arena
            self.arena = arena
            if not self.use_graph:
                self._body()
            else:
                sig = self._signature()
                if self.graph is None or moved or stale or sig != self._sig:
                    torch.cuda.current_stream().synchronize()
                    self.graph = None
                    g = torch.cuda.CUDAGraph()
                    # No cyclic garbage collection inside the capture: a dead Python cycle that owns a CUDA graph or device
                    # memory (an old decode step and its cache group reference each other) would be finalised by whatever
                    # allocation happens to trigger the collector, and a cudaFree / cudaGraphExecDestroy on the capturing
                    # thread invalidates the capture.  Collect first (captures are rare: one per pooled cache group),
                    # then keep the collector off until the capture has ended.
                    gc_was_on = gc.isenabled()
                    gc.collect()
                    gc.disable()
                    try:
                        with torch.cuda.graph(g, capture_error_mode="thread_local"):
                            self._body()
                    finally:
                        if gc_was_on:
                            gc.enable()
                    self.graph, self._sig = g, sig
                    # capture does not execute: replay below performs the step
                self.graph.replay()
            for c in self.cache:
                c.advance_host(1)


DecodeSession = _DecodeStep      # public name: one replayable decode step over a model + cache group


def _decode_step_for(model: DecoderModel, cache: List[BatchedKVCache], B: int, graph_greedy: bool) -> _DecodeStep:
    """The decode step (static buffers + captured graph) lives with the cache group it was captured over: a pooled cache
    that comes back with the same buffers replays the same graph instead of paying a ~350-node capture per request.
    ``_DecodeStep.run`` re-captures by itself when a buffer moved."""
    holder = cache[0]
    steps = holder.__dict__.setdefault("_decode_steps", {})
    key = (id(model), B, graph_greedy, len(cache))
    st = steps.get(key)
    if st is None or st.model is not model or st.cache is not cache:
        if len(steps) > 4:
            steps.clear()
        st = steps[key] = _DecodeStep(model, cache, B, graph_greedy)
    return st


def generate_step(
    prompts,
    model: DecoderModel,
    temp: float = 0.0,
    repetition_penalty: Optional[float] = None,
    repetition_context_size: Optional[int] = 20,
    top_p: float = 1.0,
    logit_bias: Optional[Dict[int, float]] = None,
    cache: Optional[List[BatchedKVCache]] = None,
) -> Generator[Tuple[torch.Tensor, torch.Tensor], None, None]:
    """Generator yielding (tokens [B,1] int64, probs [B,1] float32) forever (utils.py:315-427).

    Same contract as the reference: prefill on the whole (left-padded) prompt batch, greedy if temp == 0
    else top-p / categorical sampling, repetition_penalty raises NotImplementedError (utils.py:366-367),
    and the step for token n+1 is already in flight when token n is yielded (utils.py:420-427).
    ``probs`` is per row (the reference gathers row 0 only)."""
    if repetition_penalty:
        raise NotImplementedError("repetition_penalty not supported.")
    dev = model.device
    y = torch.as_tensor(np.asarray(prompts) if not isinstance(prompts, torch.Tensor) else prompts).to(dev)
    if y.dim() == 1:
        y = y[None, :]
    B = y.shape[0]
    kv_heads = [l.self_attn.n_kv_heads for l in model.layers]
    if cache is None:
        cache = _kv_pool.get(model.head_dim, kv_heads, B, paged=True, device=dev, owner=id(model))   # utils.py:390-394
    elif cache and cache[0]._lengths is None:
        bind_cache_group(cache, dev)
    greedy = temp == 0
    bias_idx = bias_val = None
    if logit_bias:
        bias_idx = torch.tensor(list(logit_bias.keys()), dtype=torch.int64, device=dev)
        bias_val = torch.tensor(list(logit_bias.values()), dtype=torch.float32, device=dev)
    step = _decode_step_for(model, cache, B, greedy and not logit_bias)

    def sample(logits: torch.Tensor) -> None:
        """utils.py:345-364; writes step.tok / step.probs."""
        with model._lock:
            if bias_idx is not None:
                logits[:, bias_idx] += bias_val
            if greedy:
                model.greedy_tokens(logits, step.tok, step.probs)
            else:
                p = top_p if (0 < top_p < 1.0) else 1.0
                model.sample_tokens(logits, temp, p, sample_utils.current_seed(), sample_utils.next_offset(),
                                    step.tok, step.probs)

    def check_exchange():
        """Tensor parallelism: a peer that never arrived makes the exchange kernel time out, set a status word and carry
        on with incomplete sums -- plausible but wrong tokens.  Read the word at generation boundaries (one 4-byte D2H) and
        fail loudly instead."""
        fx = getattr(model, "_tp_fused", None)
        if fx is not None and fx.status() != 0:
            raise RuntimeError("tensor-parallel exchange timed out waiting for a peer rank (status word "
                               f"{fx.status()}): results of this generation are invalid")
        # same boundary, same reason: a row appended beyond its cache capacity is dropped by the kernel (it cannot happen
        # through prepare_cache, which grows the buffers first -- this is the loud failure if that invariant ever breaks)
        dropped = ops.kv_overflow_count(reset=True)
        if dropped:
            raise RuntimeError(f"KV cache overflow: {dropped} (token, head) appends were dropped by b200p_rope_kv_append; "
                               "results of this generation are invalid")

    # prefill (eager, chunked) + first sample
    logits = model(y, cache=cache, last_only=True)[:, -1, :]
    sample(logits)
    check_exchange()
    out_y = step.tok.to(torch.int64).view(B, 1)
    out_p = step.probs.clone().view(B, 1)
    n_steps = 0
    while True:
        n_steps += 1
        if n_steps % 256 == 0:
            check_exchange()
        step.run()                                   # launches step n+1 (utils.py:422-423)
        if not step.greedy:
            sample(step.logits)
        next_y = step.tok.to(torch.int64).view(B, 1)
        next_p = step.probs.clone().view(B, 1)
        yield out_y, out_p
        out_y, out_p = next_y, next_p


# --------------------------------------------------------------------------------- text front-ends
def stream_generate(model, tokenizer, prompt: str, max_tokens: int = 100, **kwargs):
    """utils.py:429-471."""
    if not isinstance(tokenizer, TokenizerWrapper):
        tokenizer = TokenizerWrapper(tokenizer)
    prompt_tokens = torch.tensor(tokenizer.encode(prompt))[None, :]
    detokenizer = tokenizer.detokenizer
    detokenizer.reset()
    kwargs = _presize(kwargs, model, prompt_tokens, max_tokens)
    for (token, prob), n in zip(generate_step(prompt_tokens, model, **kwargs), range(max_tokens)):
        token_item = token.item()
        if token_item == tokenizer.eos_token_id:
            break
        detokenizer.add_token(token_item)
        yield detokenizer.last_segment
    detokenizer.finalize()
    yield detokenizer.last_segment


def generate(model, tokenizer, prompt: str, max_tokens: int = 100, verbose: bool = False,
             formatter: Optional[Callable] = None, **kwargs) -> str:
    """utils.py:545-616."""
    if not isinstance(tokenizer, TokenizerWrapper):
        tokenizer = TokenizerWrapper(tokenizer)
    if verbose:
        print("=" * 10)
        print("Prompt:", prompt)
    prompt_tokens = torch.tensor(tokenizer.encode(prompt))[None]
    detokenizer = tokenizer.detokenizer
    tic = time.perf_counter()
    detokenizer.reset()
    n = -1
    prompt_time = 0.0
    kwargs = _presize(kwargs, model, prompt_tokens, max_tokens)
    for (token, prob), n in zip(generate_step(prompt_tokens, model, **kwargs), range(max_tokens)):
        if n == 0:
            prompt_time = time.perf_counter() - tic
            tic = time.perf_counter()
        if token.item() == tokenizer.eos_token_id:
            break
        detokenizer.add_token(token.item())
        if verbose:
            if formatter:
                detokenizer.finalize()
                formatter(detokenizer.last_segment, prob.item())
            else:
                print(detokenizer.last_segment, end="", flush=True)
    token_count = n + 1
    detokenizer.finalize()
    if verbose:
        gen_time = time.perf_counter() - tic
        print(detokenizer.last_segment, flush=True)
        print("=" * 10)
        if token_count == 0:
            print("No tokens generated for this prompt")
            return
        print(f"Prompt: {prompt_tokens.numel() / max(prompt_time, 1e-9):.3f} tokens-per-sec")
        print(f"Generation: {(token_count - 1) / max(gen_time, 1e-9):.3f} tokens-per-sec")
    return detokenizer.text


def batch_generate(model, tokenizer, prompts: List[str], max_tokens: int = 100, verbose: bool = False,
                   format_prompts: bool = True, formatter: Optional[Callable] = None, **kwargs) -> List[str]:
    """utils.py:473-543: chat-template, left-pad (pad = eos if unset), run max_tokens steps with no early
    stop, decode and cut at the eos / pad strings."""
    if not isinstance(tokenizer, TokenizerWrapper):
        tokenizer = TokenizerWrapper(tokenizer)
    if verbose:
        print("=" * 10)
    if format_prompts:
        prompts_fm = [[{"role": "user", "content": p}] for p in prompts]
        prompts_fm = [tokenizer.apply_chat_template(p, add_generation_prompt=True, tokenize=False) for p in prompts_fm]
    else:
        prompts_fm = prompts
    tokenizer._tokenizer.padding_side = "left"
    if tokenizer.pad_token is None:
        tokenizer._tokenizer.pad_token = tokenizer.eos_token
        tokenizer._tokenizer.pad_token_id = tokenizer.eos_token_id
    prompts_toks = torch.tensor(tokenizer._tokenizer(prompts_fm, padding=True)["input_ids"])
    tic = time.perf_counter()
    output_toks = []
    prompt_time = 0.0
    kwargs = _presize(kwargs, model, prompts_toks, max_tokens)
    for (tokens, _), n in zip(generate_step(prompts_toks, model, **kwargs), range(max_tokens)):
        if n == 0:
            tokens[0, 0].item()                      # first token materialised = prefill done
            prompt_time = time.perf_counter() - tic
            tic = time.perf_counter()
        output_toks.append(tokens)
    output_toks = torch.cat(output_toks, dim=1)
    out_list = output_toks.tolist()                  # one D2H for the whole generation
    responses = [r.split(tokenizer.eos_token)[0].split(tokenizer.pad_token)[0] for r in tokenizer.batch_decode(out_list)]
    if verbose:
        gen_time = time.perf_counter() - tic
        print(f"Prompt: {prompts_toks.numel() / max(prompt_time, 1e-9):.3f} tokens-per-sec")
        print(f"Generation: {output_toks.numel() / max(gen_time, 1e-9):.3f} tokens-per-sec")
        for prompt, response in zip(prompts, responses):
            print("=" * 10)
            print("Prompt:", prompt)
            print(response)
    return responses


def batch_stream_generate_text(model, tokenizer, prompts_tokens, max_tokens: int, **kwargs):
    """utils.py:983-1075: yields, per step, a list of (text_delta | None, finish_reason | None) per row."""
    prompts_tokens = torch.as_tensor(np.asarray(prompts_tokens) if not isinstance(prompts_tokens, torch.Tensor) else prompts_tokens)
    batch_size = prompts_tokens.shape[0]
    if not isinstance(tokenizer, TokenizerWrapper):
        tokenizer = TokenizerWrapper(tokenizer)
    detokenizers = [copy.deepcopy(tokenizer.detokenizer) for _ in range(batch_size)]
    for d in detokenizers:
        d.reset()
    active = [True] * batch_size
    counts = [0] * batch_size
    eos_token_id = tokenizer.eos_token_id
    gs_kwargs = _presize({k: v for k, v in kwargs.items() if k != "repetition_penalty"}, model, prompts_tokens, max_tokens)
    for (batch_next, _), _ in zip(generate_step(prompts_tokens, model, **gs_kwargs), range(max_tokens)):
        ids = batch_next.view(-1).tolist()           # ONE D2H per step (the reference does B .item() syncs)
        deltas: List[Tuple[Optional[str], Optional[str]]] = [(None, None)] * batch_size
        any_active = False
        for i in range(batch_size):
            if not active[i]:
                continue
            any_active = True
            token_id = ids[i]
            counts[i] += 1
            delta, reason = None, None
            if token_id == eos_token_id:
                active[i] = False
                detokenizers[i].finalize()
                delta = detokenizers[i].last_segment
                reason = "stop"
            else:
                detokenizers[i].add_token(token_id)
                delta = detokenizers[i].last_segment
            if active[i] and counts[i] >= max_tokens:
                active[i] = False
                if not reason:
                    detokenizers[i].finalize()
                    final_segment = detokenizers[i].last_segment
                    if final_segment:
                        delta = final_segment
                    reason = "length"
            deltas[i] = (delta, reason)
        yield deltas
        if not any_active:
            break


async def batch_generate_text(model, tokenizer, prompts: List[str], max_tokens: int = 100, temp: float = 0.7,
                              top_p: float = 1.0, disable_prefix_cache: bool = False,
                              max_context_length: Optional[int] = None) -> List[Tuple[str, int, int]]:
    """utils.py:1087-1346: tokenise left-padded, split off the longest common prefix, prefill it once
    for the whole batch, run the suffix batch through generate_step with per-row EOS / max_tokens
    bookkeeping, return (text, n_prompt_tokens, n_completion_tokens) per prompt."""
    if not prompts:
        return []
    loop = asyncio.get_running_loop()
    tk = tokenizer._tokenizer
    if tk.pad_token_id is None:
        tk.pad_token_id = tk.eos_token_id if tk.eos_token_id is not None else 0
    original_side = tk.padding_side
    tk.padding_side = "left"
    if max_context_length is not None:
        eff = max_context_length
    else:
        eff = 2048
        cand = getattr(tk, "model_max_length", None)
        try:
            if cand is not None and int(cand) > 0:
                eff = min(int(cand), 65536)
        except (ValueError, TypeError):
            pass
    try:
        batch = tk(prompts, return_tensors="np", padding="longest", truncation=True, max_length=eff)
    finally:
        tk.padding_side = original_side
    prompt_np = batch["input_ids"].astype(np.int64)
    mask_np = batch["attention_mask"]
    n_prompt = [int(np.sum(m)) for m in mask_np]

    def _synchronous_generation():
        B = prompt_np.shape[0]
        eos = tokenizer.eos_token_id
        if isinstance(eos, set):
            eos = next(iter(eos)) if eos else None
        seqs = [ids[m.astype(bool)].tolist() for ids, m in zip(prompt_np, mask_np)]
        lcp = 0
        if len(seqs) > 1:
            for pos in range(min(len(s) for s in seqs)):
                if all(s[pos] == seqs[0][pos] for s in seqs):
                    lcp += 1
                else:
                    break
        elif len(seqs) == 1:
            lcp = len(seqs[0])
        suffixes = [s[lcp:] for s in seqs]
        est = max((len(s) for s in suffixes), default=0)
        desired = max(256, min(8192, int(((est + max_tokens) // 256 + 1) * 256)))       # utils.py:1234
        kv_heads = [l.self_attn.n_kv_heads for l in model.layers]
        caches = _kv_pool.get(model.head_dim, kv_heads, B, step=desired, paged=True, device=model.device, owner=id(model))
        if lcp > 0 and not disable_prefix_cache:
            lcp_ids = torch.tensor(np.tile(np.array(seqs[0][:lcp], dtype=np.int64), (B, 1)))
            model(lcp_ids, cache=caches, last_only=True)                                 # utils.py:1254-1259
        elif lcp > 0:
            suffixes = seqs
        mx_len = max((len(s) for s in suffixes), default=0)
        pad_id = eos if eos is not None else 0
        if mx_len > 0:
            suffix_batch = np.array([[pad_id] * (mx_len - len(s)) + s for s in suffixes], dtype=np.int64)
        else:
            suffix_batch = np.array([[pad_id]] * B, dtype=np.int64)
        generated = [[] for _ in range(B)]
        active = [True] * B
        counts = [0] * B
        for step_num, (nxt, _) in enumerate(generate_step(suffix_batch, model, cache=caches, temp=temp, top_p=top_p)):
            if step_num >= max_tokens:
                break
            ids = nxt.view(-1).tolist()
            any_active = False
            for i in range(B):
                if not active[i]:
                    continue
                any_active = True
                if ids[i] == eos or counts[i] >= max_tokens:
                    active[i] = False
                else:
                    generated[i].append(ids[i])
                    counts[i] += 1
            if not any_active:
                break
        return [(tokenizer.decode(generated[i], skip_special_tokens=True), n_prompt[i], len(generated[i])) for i in range(B)]

    return await loop.run_in_executor(None, _synchronous_generation)


batch_generate_text_util = batch_generate_text


# --------------------------------------------------------------------------------- loading
def load_config(model_path: Path) -> dict:
    try:
        with open(Path(model_path) / "config.json", "r") as f:
            return json.load(f)
    except FileNotFoundError:
        logging.error(f"Config file not found in {model_path}")
        raise


def load_model(model_path: Path, lazy: bool = False, model_config: dict = {}, device="cuda",
               tp_rank: int = 0, tp_world: int = 1) -> DecoderModel:
    """utils.py:630-708: config.json -> arch classes -> safetensors (MLX layout) -> quantized modules.
    Only tensors that exist as module parameters are taken (utils.py:693-699).  With tp_world > 1 each
    rank keeps its column / row shard (tp.shard_weights)."""
    model_path = Path(model_path)
    config = load_config(model_path)
    config.update(model_config)
    weight_files = glob.glob(str(model_path / "model*.safetensors")) or glob.glob(str(model_path / "weight*.safetensors"))
    if not weight_files:
        logging.error(f"No safetensors found in {model_path}")
        raise FileNotFoundError(f"No safetensors found in {model_path}")
    from safetensors import safe_open
    weights: Dict[str, torch.Tensor] = {}
    for wf in weight_files:
        with safe_open(wf, framework="pt", device="cpu") as f:
            for k in f.keys():
                weights[k] = f.get_tensor(k)
    model_class, model_args_class = _get_classes(config)
    if "quantization" not in config:
        raise ValueError("only MLX-quantized checkpoints (config['quantization']) are on this build's hot path")
    model = model_class(model_args_class.from_dict(config))
    if hasattr(model, "sanitize"):
        weights = model.sanitize(weights)
    known = set(model.parameter_names())
    dropped = [k for k in weights if k not in known]
    if dropped:
        logging.warning(f"Filtering out {len(dropped)} unmatched weight tensors during load.")
    weights = {k: v for k, v in weights.items() if k in known}
    if tp_world > 1:
        from . import tp
        weights = tp.shard_weights(weights, config, tp_rank, tp_world)
    model.load_weights(weights, device=device)
    return model


# --------------------------------------------------------------------------------- quantiser / converter (row f3)
QUANTIZED_LEAVES = ("q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj", "embed_tokens", "lm_head")


def quantize_model(weights: Dict[str, torch.Tensor], config: dict, q_group_size: int = 64, q_bits: int = 4,
                   device="cuda") -> Tuple[Dict[str, torch.Tensor], dict]:
    """utils.py:886-905 on a dense state dict: every ``nn.Linear`` / ``nn.Embedding`` weight (nn.quantize replaces all
    modules that have ``to_quantized``) becomes ``weight`` uint32 codes + ``scales`` + ``biases`` through b200p_quantize
    (mx.quantize, bit-exact against the oracle); norms pass through.  Returns (weights, config with "quantization")."""
    out: Dict[str, torch.Tensor] = {}
    for name, t in weights.items():
        mod, _, leaf = name.rpartition(".")
        if leaf == "weight" and t.dim() == 2 and mod.rsplit(".", 1)[-1] in QUANTIZED_LEAVES and t.shape[1] % q_group_size == 0:
            wq, sc, bi = ops.quantize(t.to(device).contiguous(), q_group_size, q_bits)
            out[name], out[mod + ".scales"], out[mod + ".biases"] = wq.cpu(), sc.cpu(), bi.cpu()
        else:
            out[name] = t
    qcfg = copy.deepcopy(config)
    qcfg["quantization"] = {"group_size": q_group_size, "bits": q_bits}
    return out, qcfg


def convert(hf_path: str, mlx_path: str = "mlx_model", quantize: bool = False, q_group_size: int = 64, q_bits: int = 4,
            dtype: str = "float16", upload_repo: Optional[str] = None, revision: Optional[str] = None,
            dequantize: bool = False):
    """utils.py:934-981 for local checkpoints: HF dense safetensors -> (quantised) MLX-format checkpoint
    (``model.safetensors`` with metadata format=mlx + ``config.json`` with the quantization block)."""
    if quantize and dequantize:
        raise ValueError("Choose either quantize or dequantize, not both.")
    if dequantize or upload_repo is not None:
        raise NotImplementedError("dequantize / upload are outside this build (no network; decode path only)")
    from safetensors import safe_open
    from . import synth
    model_path = get_model_path(hf_path, revision)
    config = load_config(model_path)
    files = glob.glob(str(model_path / "model*.safetensors")) or glob.glob(str(model_path / "*.safetensors"))
    if not files:
        raise FileNotFoundError(f"No safetensors found in {model_path}")
    tdt = torch.float16 if quantize else getattr(torch, dtype)
    weights: Dict[str, torch.Tensor] = {}
    for wf in files:
        with safe_open(wf, framework="pt", device="cpu") as f:
            for k in f.keys():
                t = f.get_tensor(k)
                weights[k] = t.to(tdt) if t.is_floating_point() else t
    if quantize:
        weights, config = quantize_model(weights, config, q_group_size, q_bits)
    synth.save_checkpoint(str(mlx_path), config, weights)
    return Path(mlx_path)


def load_adapters(model: DecoderModel, adapter_path) -> DecoderModel:
    """mlx_lm.tuner.utils.load_adapters as the reference calls it (utils.py:737-741): ``adapter_config.json``
    {fine_tune_type, num_layers, lora_parameters{rank, scale, dropout, keys}} + ``adapters.safetensors`` (lora_init.py:140-149)."""
    from safetensors import safe_open
    adapter_path = Path(adapter_path)
    cfg_file = adapter_path / "adapter_config.json"
    if not cfg_file.exists():
        raise FileNotFoundError(f"The adapter path does not exist: {adapter_path}")
    cfg = json.load(open(cfg_file))
    if cfg.get("fine_tune_type", "lora") != "lora":
        raise NotImplementedError(f"fine_tune_type {cfg.get('fine_tune_type')} is not built (LoRA only)")
    lp = cfg.get("lora_parameters", {})
    weights: Dict[str, torch.Tensor] = {}
    with safe_open(str(adapter_path / "adapters.safetensors"), framework="pt", device="cpu") as f:
        for k in f.keys():
            weights[k] = f.get_tensor(k)
    model.apply_lora(weights, lp.get("scale", 10.0), tuple(lp.get("keys", ("self_attn.q_proj", "self_attn.v_proj"))))
    return model


def load(path_or_hf_repo: str, tokenizer_config={}, model_config={}, adapter_path: Optional[str] = None,
         lazy: bool = False):
    """utils.py:711-747 -> (model, TokenizerWrapper)."""
    model_path = get_model_path(path_or_hf_repo)
    model = load_model(model_path, lazy, model_config)
    if adapter_path is not None:
        load_adapters(model, adapter_path)
    tokenizer = load_tokenizer(model_path, tokenizer_config)
    return model, tokenizer
