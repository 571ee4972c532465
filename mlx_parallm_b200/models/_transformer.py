"""Shared decoder implementation behind models/llama.py and models/qwen3.py.

Same module tree, parameter names and call signatures as the reference graphs
(mlx_parallm/models/llama.py:49-271, mlx_parallm/models/qwen3.py:23-226), re-hosted on the
sm_100a kernels of libb200parallm.so:

    embed (gather+dequant)                                   b200p_embed
    per layer:  RMSNorm                                      b200p_rmsnorm
                fused q|k|v projection                       b200p_qmm (EPI_STORE)
                (q/k head norm) + RoPE + KV append           b200p_rope_kv_append
                GQA attention over the cache                 b200p_attn
                o_proj + residual                            b200p_qmm (EPI_ADD_RESIDUAL)
                RMSNorm                                      b200p_rmsnorm
                fused gate|up projection + silu*mul          b200p_qmm (EPI_SILU_MUL)
                down_proj + residual                         b200p_qmm (EPI_ADD_RESIDUAL)
    final RMSNorm + lm_head (float32 logits)                 b200p_rmsnorm, b200p_qmm (EPI_STORE_F32)

The residual stream is float32 (the reference's stream is float32 from layer 0 on because
PagedKVCache promotes); GEMM inputs and the KV cache are the model's 16-bit storage
type.  Tensor parallelism (column-split q/k/v/gate/up, row-split o/down, vocab-split head) is
applied at load time by ``tp.shard_*`` and adds one all-reduce after each row-split projection.
"""
from __future__ import annotations

import threading
from typing import Dict, List, Optional

import torch

from .. import ops
from .._lib import EPI_ADD_RESIDUAL, EPI_SILU_MUL, EPI_STORE, EPI_STORE_F32
from .base import BatchedKVCache, PagedKVCache

MAX_TOKENS_PER_FORWARD = 8192     # chunk size (B*L) of one prefill pass through the layers
# Up to this many rows the fused q|k|v projection accumulates in an fp32 buffer through the stream-K kernel (csrc/qmm_sk.cu:
# partial tiles summed by TMA reduce-adds, no tail; rope_kv_append rounds to the storage type and re-zeroes the buffer).
# q|k|v has 24 output tiles for 148 SMs, so it must split K; measured at 64 rows (profiles/r2_run11_*): 16.8 us through the
# whole-tile kernel's workspace split against 8-10 us this way.  Above it (prefill) whole 16-bit tiles are stored directly.
ACC32_MAX_ROWS = 128
# gate|up: one GPU's full matrix (117 M weights on Llama-3-8B) goes through the whole-tile kernel with the fused silu*mul
# epilogue (27.9 us at 64 rows; stream-K into fp32 + the silu_mul kernel: 22.7 + 5.9 us, and 22 MB of fp32 round trip).  A
# tensor-parallel SHARD of it has too few rows to fill 148 SMs with whole tiles and is latency-bound there, so up to this
# many weights it takes the stream-K accumulate path + silu_mul (profiles/r2_run17_*, r2_run18_*, 64 rows: TP=2 shard
# 14.9 + 4.2 vs 23.5 us, TP=4 10.8 + 3.5 vs 17.3, TP=8 7.4 + 3.1 vs 13.7).
GATE_UP_ACC32_MAX_WEIGHTS = 100 * 1000 * 1000
USE_PERM8 = True                  # 4-bit 16-bit models: GEMM inputs in the PERM8 column order (include/b200p.h)
# |scale| bound under which an f16 matrix may take the subnormal-operand dequant (B200P_W_SMALLSCALE, csrc/dequant.cuh):
# the kernel needs s * 2^20 finite in half precision, i.e. |s| < 0.0625; checked once per matrix at load time.
SMALL_SCALE_BOUND = 0.06


class QuantizedLinear:
    """nn.QuantizedLinear stand-in: weight uint32 [N, K*bits/32] (held as int32), scales/biases [N, K/gs]."""

    def __init__(self, in_features: int, out_features: int, group_size: int = 64, bits: int = 4):
        self.in_features, self.out_features = in_features, out_features
        self.group_size, self.bits = group_size, bits
        self.weight = self.scales = self.biases = None
        self.small_scale = False

    def set(self, weight, scales, biases, small_scale: Optional[bool] = None):
        self.weight, self.scales, self.biases = weight, scales, biases
        self.out_features = weight.shape[0]
        self.in_features = weight.shape[1] * 32 // self.bits
        if small_scale is None:      # one reduction + host read per matrix, at load time only
            small_scale = (scales.dtype == torch.float16 and scales.is_cuda and scales.numel() > 0
                           and float(scales.abs().max()) < SMALL_SCALE_BOUND)
        self.small_scale = bool(small_scale)

    def __call__(self, x, epilogue=EPI_STORE, out=None, workspace=None, x_perm=False, y_perm=False):
        lead = x.shape[:-1]
        y = ops.qmm(x.reshape(-1, x.shape[-1]), self.weight, self.scales, self.biases, self.group_size,
                    self.bits, epilogue, out, workspace, x_perm=x_perm, y_perm=y_perm, small_scale=self.small_scale)
        return y if out is not None else y.reshape(*lead, -1)


class QuantizedEmbedding(QuantizedLinear):
    """nn.QuantizedEmbedding stand-in (llama.py:212; .as_linear for tied heads, llama.py:250)."""

    def embed(self, ids_i32, out=None):
        return ops.embed(ids_i32, self.weight, self.scales, self.biases, self.group_size, self.bits, out=out)

    def as_linear(self, x, **kw):
        return QuantizedLinear.__call__(self, x, **kw)


class RMSNorm:
    def __init__(self, dims: int, eps: float):
        self.weight = None
        self.eps = eps
        self.dims = dims


class Attention:
    def __init__(self, args, head_dim: int, qk_norm: bool):
        self.n_heads = args.num_attention_heads
        self.n_kv_heads = args.num_key_value_heads
        self.head_dim = head_dim
        self.scale = head_dim ** -0.5
        gs, bits = args._group_size, args._bits
        h = args.hidden_size
        self.q_proj = QuantizedLinear(h, self.n_heads * head_dim, gs, bits)
        self.k_proj = QuantizedLinear(h, self.n_kv_heads * head_dim, gs, bits)
        self.v_proj = QuantizedLinear(h, self.n_kv_heads * head_dim, gs, bits)
        self.o_proj = QuantizedLinear(self.n_heads * head_dim, h, gs, bits)
        self.qkv_proj = QuantizedLinear(h, (self.n_heads + 2 * self.n_kv_heads) * head_dim, gs, bits)
        self.lora: Dict[str, tuple] = {}      # "q_proj" / "v_proj" -> (At [16,K], Bt [N,16], scale): set by apply_lora
        self.q_norm = RMSNorm(head_dim, args.rms_norm_eps) if qk_norm else None
        self.k_norm = RMSNorm(head_dim, args.rms_norm_eps) if qk_norm else None


class MLP:
    def __init__(self, args):
        gs, bits = args._group_size, args._bits
        h, inter = args.hidden_size, args.intermediate_size
        self.gate_proj = QuantizedLinear(h, inter, gs, bits)
        self.up_proj = QuantizedLinear(h, inter, gs, bits)
        self.down_proj = QuantizedLinear(inter, h, gs, bits)
        self.gate_up_proj = QuantizedLinear(h, 2 * inter, gs, bits)


class TransformerBlock:
    def __init__(self, args, head_dim: int, qk_norm: bool):
        self.self_attn = Attention(args, head_dim, qk_norm)
        self.mlp = MLP(args)
        self.input_layernorm = RMSNorm(args.hidden_size, args.rms_norm_eps)
        self.post_attention_layernorm = RMSNorm(args.hidden_size, args.rms_norm_eps)
        self.args = args


class DecoderStack:
    """LlamaModel / Qwen3Model (llama.py:194-231, qwen3.py:146-185)."""

    def __init__(self, args, head_dim: int, qk_norm: bool):
        self.args = args
        self.vocab_size = args.vocab_size
        self.num_hidden_layers = args.num_hidden_layers
        assert self.vocab_size > 0
        self.embed_tokens = QuantizedEmbedding(args.hidden_size, args.vocab_size, args._group_size, args._bits)
        self.layers = [TransformerBlock(args, head_dim, qk_norm) for _ in range(args.num_hidden_layers)]
        self.norm = RMSNorm(args.hidden_size, args.rms_norm_eps)


def _fuse_rows(parts: List[QuantizedLinear], fused: QuantizedLinear):
    """Concatenate row blocks into one MLX-layout matrix and re-point the parts at views of it."""
    w = torch.cat([p.weight for p in parts], 0).contiguous()
    s = torch.cat([p.scales for p in parts], 0).contiguous()
    b = torch.cat([p.biases for p in parts], 0).contiguous()
    flags = [p.small_scale for p in parts]
    fused.set(w, s, b, small_scale=all(flags))
    r = 0
    for p, fl in zip(parts, flags):
        n = p.weight.shape[0]
        p.set(w[r:r + n], s[r:r + n], b[r:r + n], small_scale=fl)
        r += n


class DecoderModel:
    """``Model`` of the reference: callable(inputs[B,L], cache=None) -> logits [B, L, V] (float32)."""

    qk_norm = False

    def __init__(self, args):
        self.args = args
        self.model_type = args.model_type
        q = getattr(args, "quantization", None) or {}
        args._group_size = q.get("group_size", 64)
        args._bits = q.get("bits", 4)
        hd = self._head_dim(args)
        self._hd = hd
        self.model = DecoderStack(args, hd, self.qk_norm)
        self.lm_head = None
        if not args.tie_word_embeddings:
            self.lm_head = QuantizedLinear(args.hidden_size, args.vocab_size, args._group_size, args._bits)
        self.dtype: Optional[torch.dtype] = None
        self.device: Optional[torch.device] = None
        self._arenas: Dict[int, dict] = {}
        # One forward at a time per model instance (SURVEY 8b "Threading"): the scratch arenas, the self-cleaning qmm
        # workspace and the fp32 accumulators are shared by every caller of this model, and the reference server runs
        # generate_step from an executor thread and from the event loop at once (utils.py:1345 vs server/main.py:1358).
        # Held around every eager forward, every decode-step replay and every graph capture.
        self._lock = threading.RLock()
        self.tp_rank, self.tp_world, self.tp_group = 0, 1, None
        self._allreduce = None
        rs = getattr(args, "rope_scaling", None)
        self.rope_scale = 1.0
        if rs and (rs.get("type") == "linear" or rs.get("rope_type") == "linear"):
            self.rope_scale = 1.0 / rs["factor"]                        # llama.py:69-76

    # -- reference properties ----------------------------------------------------------------
    @staticmethod
    def _head_dim(args) -> int:
        return args.hidden_size // args.num_attention_heads              # llama.py:55

    @property
    def layers(self):
        return self.model.layers

    @property
    def head_dim(self):
        return self._hd

    @property
    def n_kv_heads(self):
        """KV heads held by THIS rank (what a cache for this model instance must be built with)."""
        return self.model.layers[0].self_attn.n_kv_heads

    def sanitize(self, weights):
        return {k: v for k, v in weights.items() if "self_attn.rotary_emb.inv_freq" not in k}   # llama.py:254-258

    def eval(self):
        return self

    # -- parameters ----------------------------------------------------------------------------
    def parameter_names(self) -> List[str]:
        names = []
        for pre in ["model.embed_tokens"] + (["lm_head"] if self.lm_head is not None else []):
            names += [f"{pre}.{s}" for s in ("weight", "scales", "biases")]
        names.append("model.norm.weight")
        for i in range(len(self.model.layers)):
            p = f"model.layers.{i}."
            for proj in ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj",
                         "mlp.gate_proj", "mlp.up_proj", "mlp.down_proj"):
                names += [f"{p}{proj}.{s}" for s in ("weight", "scales", "biases")]
            names += [p + "input_layernorm.weight", p + "post_attention_layernorm.weight"]
            if self.qk_norm:
                names += [p + "self_attn.q_norm.weight", p + "self_attn.k_norm.weight"]
        return names

    def _module(self, path: str):
        obj = self
        for part in path.split("."):
            obj = obj[int(part)] if part.isdigit() else getattr(obj, part)
        return obj

    def load_weights(self, weights: Dict[str, torch.Tensor], device="cuda", strict: bool = True):
        """Take MLX-format tensors (utils.py:654-708 key naming), move them to the device, fuse q|k|v and
        gate|up row blocks.  Under TP the tensors must already be this rank's shards (tp.shard_weights)."""
        device = torch.device(device)
        missing = [n for n in self.parameter_names() if n not in weights]
        if missing and strict:
            raise KeyError(f"missing parameters: {missing[:4]}{'...' if len(missing) > 4 else ''}")
        dts = {weights[n].dtype for n in self.parameter_names() if n in weights and n.endswith(".scales")}
        if len(dts) != 1:
            raise TypeError(f"scales must share one dtype, got {dts}")
        self.dtype = dts.pop()
        self.device = device
        a0 = self.args
        self._fast16 = self.dtype in (torch.float16, torch.bfloat16)
        # PERM8 needs every GEMM input row to be made of aligned groups of 8 (hidden, heads*head_dim, intermediate)
        self._perm8 = (self._fast16 and a0._bits == 4 and a0.hidden_size % 8 == 0 and self._hd % 8 == 0
                       and a0.intermediate_size % 8 == 0 and USE_PERM8)

        def put(t):
            if t.dtype == torch.uint32:
                t = t.view(torch.int32)
            return t.to(device).contiguous()

        groups: Dict[str, dict] = {}
        for name in self.parameter_names():
            if name not in weights:
                continue
            mod, leaf = name.rsplit(".", 1)
            groups.setdefault(mod, {})[leaf] = put(weights[name])
        for mod, g in groups.items():
            m = self._module(mod)
            if isinstance(m, QuantizedLinear):
                m.set(g["weight"], g["scales"], g["biases"])
            else:
                m.weight = g["weight"].to(self.dtype)
        for layer in self.model.layers:
            a, f = layer.self_attn, layer.mlp
            a.n_heads = a.q_proj.weight.shape[0] // a.head_dim          # local (TP-sharded) head counts
            a.n_kv_heads = a.k_proj.weight.shape[0] // a.head_dim
            _fuse_rows([a.q_proj, a.k_proj, a.v_proj], a.qkv_proj)
            _fuse_rows([f.gate_proj, f.up_proj], f.gate_up_proj)
        self._arenas.clear()
        return self

    def apply_lora(self, adapters: Dict[str, torch.Tensor], scale: float, keys=("self_attn.q_proj", "self_attn.v_proj")):
        """Attach mlx_lm-format LoRA adapters (``model.layers.{i}.self_attn.q_proj.lora_a [K, r]`` / ``.lora_b [r, N]``;
        lora_init.py:63-73, 140-149) to the attention projections.  Stored transposed and rank-padded to 16 for the
        mma kernel; lora_a's k axis follows the PERM8 activation layout when the model uses it.  Pass {} to detach.

        Tensor parallelism: q_proj / v_proj are column-parallel, so ``lora_a`` (applied to the replicated input) is used
        whole and ``lora_b`` is cut to this rank's output columns -- the same row range ``tp.shard_weights`` keeps of the
        projection.  Hot swap (rl_training/weight_updater.py:66-73): the adapter set is built aside and swapped in under the
        model's forward lock, so a generation in flight sees either the old or the new set, never a mix."""
        if not self._fast16:
            raise NotImplementedError("LoRA needs a 16-bit model (b200p_lora_add)")
        new_sets = []
        for i, layer in enumerate(self.model.layers):
            new_sets.append({})
            for key in keys:
                if key not in ("self_attn.q_proj", "self_attn.v_proj"):
                    raise NotImplementedError(f"LoRA on {key} is not built (reference default: q_proj, v_proj)")
                pa, pb = f"model.layers.{i}.{key}.lora_a", f"model.layers.{i}.{key}.lora_b"
                if pa not in adapters:
                    continue
                A = adapters[pa].to(self.device, self.dtype)           # [K, r]
                Bm = adapters[pb].to(self.device, self.dtype)          # [r, N]
                n_local = getattr(layer.self_attn, key.split(".")[-1]).weight.shape[0]
                if Bm.shape[1] != n_local:                             # full-width adapter on a column shard
                    if Bm.shape[1] != n_local * self.tp_world:
                        raise ValueError(f"{pb}: {Bm.shape[1]} output columns do not match the projection "
                                         f"({n_local} x tp {self.tp_world})")
                    Bm = Bm[:, self.tp_rank * n_local:(self.tp_rank + 1) * n_local]
                K, r = A.shape
                if r > 16:
                    raise NotImplementedError(f"LoRA rank {r} > 16")
                At = torch.zeros((16, K), dtype=self.dtype, device=self.device)
                At[:r] = A.t()
                if self._perm8:                                        # x arrives as [0,4,1,5,2,6,3,7] inside each group of 8
                    At = At.view(16, K // 8, 8)[:, :, [0, 4, 1, 5, 2, 6, 3, 7]].reshape(16, K)
                Bt = torch.zeros((Bm.shape[1], 16), dtype=self.dtype, device=self.device)
                Bt[:, :r] = Bm.t()
                new_sets[i][key.split(".")[-1]] = (At.contiguous(), Bt.contiguous(), float(scale))
        with self._lock:
            for layer, st in zip(self.model.layers, new_sets):
                layer.self_attn.lora = st
            self._arenas.clear()          # captured decode graphs hold the old adapter pointers: force a re-capture
        return self

    def set_tensor_parallel(self, rank: int, world: int, allreduce):
        """allreduce(tensor_f32) sums in place across ranks on the current stream."""
        self.tp_rank, self.tp_world, self._allreduce = rank, world, allreduce

    # -- sampling heads (overridden by the vocab-parallel variant in tp.py) ------------------------
    def graph_safe(self, B: int) -> bool:
        """Every kernel of a decode step is a graph-capturable b200p launch (there is no library or eager fallback)."""
        return True

    def greedy_tokens(self, logits, tok_out, probs_out=None):
        """mx.argmax(logits, -1) (utils.py:352) into static buffers."""
        ops.argmax(logits, out=tok_out, probs_out=probs_out)

    def sample_tokens(self, logits, temp, top_p, seed, offset, tok_out, probs_out=None):
        """top_p_sampling / categorical(logits / temp) (utils.py:354-361)."""
        ops.sample_top_p(logits, temp, top_p, seed, offset, out=tok_out, probs_out=probs_out)

    # -- caches ----------------------------------------------------------------------------------
    def make_cache(self, batch_size: int, step: Optional[int] = None, paged: bool = True) -> List[BatchedKVCache]:
        klass = PagedKVCache if paged else BatchedKVCache
        caches = [klass(self.head_dim, l.self_attn.n_kv_heads, batch_size) for l in self.model.layers]
        bind_cache_group(caches, self.device, step)
        return caches

    # -- scratch ---------------------------------------------------------------------------------
    def _arena(self, M: int, B: int) -> dict:
        # keyed by the launching stream as well: callers on different streams (threads that set their own) must not share
        # scratch; callers on ONE stream are serialised by self._lock (launch order == execution order)
        key = (M, B, torch.cuda.current_stream(self.device).cuda_stream)
        a = self._arenas.get(key)
        if a is None:
            dev, dt, args = self.device, self.dtype, self.args
            att = self.model.layers[0].self_attn
            inter_local = self.model.layers[0].mlp.down_proj.in_features
            nqkv = (att.n_heads + 2 * att.n_kv_heads) * att.head_dim
            head = self.lm_head if self.lm_head is not None else self.model.embed_tokens
            dims = [(nqkv, args.hidden_size), (args.hidden_size, att.n_heads * att.head_dim),
                    (2 * inter_local, args.hidden_size), (args.hidden_size, inter_local),
                    (head.out_features, args.hidden_size)]
            ws_q = max(max(ops.qmm_workspace_bytes(M, n, k) for n, k in dims), 16)
            ws_a = max(ops.attn_workspace_bytes(B, M // B, att.n_heads, att.head_dim), 16)
            a = dict(
                h=torch.empty((M, args.hidden_size), dtype=torch.float32, device=dev),
                x=torch.empty((M, args.hidden_size), dtype=dt, device=dev),
                qkv=torch.empty((M, nqkv), dtype=dt, device=dev),
                # zero once: the consumer kernel (rope_kv_append) re-zeroes it behind its reads
                qkv32=torch.zeros((M, nqkv), dtype=torch.float32, device=dev) if self._fast16 and M <= ACC32_MAX_ROWS else None,
                gu32=(torch.zeros((M, 2 * inter_local), dtype=torch.float32, device=dev)
                      if self._fast16 and M <= ACC32_MAX_ROWS and 2 * inter_local * args.hidden_size <= GATE_UP_ACC32_MAX_WEIGHTS
                      else None),
                q=torch.empty((M, att.n_heads * att.head_dim), dtype=dt, device=dev),
                att=torch.empty((M, att.n_heads * att.head_dim), dtype=dt, device=dev),
                act=torch.empty((M, inter_local), dtype=dt, device=dev),
                xl=torch.empty((M, args.hidden_size), dtype=dt, device=dev),
                # split-K tile counters live at the head of the qmm workspace: zero once, kernels self-clean
                ws=torch.zeros((ws_q,), dtype=torch.uint8, device=dev),
                ws_attn=torch.empty((ws_a,), dtype=torch.uint8, device=dev),
            )
            fx = getattr(self, "_tp_fused", None)
            if fx is not None and fx.mode == "rows":
                a["x"] = fx.x_view(M, dt)          # filled on every rank by the row-sharded exchange (tp.PeerExchange)
            if len(self._arenas) > 8:
                self._arenas.clear()
            self._arenas[key] = a
        return a

    # -- forward ---------------------------------------------------------------------------------
    def _layers_forward(self, ids_i32: torch.Tensor, B: int, L: int, cache: List[BatchedKVCache], a: dict):
        """Run embed + all blocks for B*L tokens whose rows sit at cache lengths + [0, L)."""
        args = self.args
        self.model.embed_tokens.embed(ids_i32, out=a["h"])
        tp = self.tp_world > 1
        # 16-bit models with 4-bit weights: GEMM inputs travel in the PERM8 column order the tensor path unpacks nibbles
        # in.  gate|up leaves its GEMM as silu(gate)*up in the storage type (whole output tiles per SM, fused epilogue);
        # q|k|v accumulates in fp32 at decode sizes (ACC32_MAX_ROWS) and is rounded by rope_kv_append -- one rounding of
        # the fp32 sum either way, which is what nn.QuantizedLinear returns (llama.py:93).
        pm = self._perm8
        fx = getattr(self, "_tp_fused", None) if tp else None      # peer-memory exchange (tp.PeerExchange) or NCCL
        layers = self.model.layers
        M = a["h"].shape[0]
        rows_mode = fx is not None and fx.mode == "rows"
        a["x_final"] = False
        for li, (layer, c) in enumerate(zip(layers, cache)):
            at, mlp = layer.self_attn, layer.mlp
            lengths = c.lengths(self.device)
            if fx is None or li == 0:      # with the fused exchange the previous layer's down_proj exchange wrote x
                ops.rmsnorm(a["h"], layer.input_layernorm.weight, layer.input_layernorm.eps, out=a["x"], out_perm=pm)
            acc32 = a["qkv32"] is not None                 # 16-bit model at decode sizes: fp32 accumulator (zero on entry)
            qkv_buf = a["qkv32"] if acc32 else a["qkv"]
            at.qkv_proj(a["x"], EPI_ADD_RESIDUAL if acc32 else EPI_STORE, out=qkv_buf, workspace=a["ws"], x_perm=pm)
            if at.lora:                    # LoRA on q_proj / v_proj (A.9): low-rank update of the q / v column blocks
                for name, (At, Bt, scale) in at.lora.items():
                    col0 = 0 if name == "q_proj" else (at.n_heads + at.n_kv_heads) * at.head_dim
                    ops.lora_add(a["x"], At, Bt, qkv_buf, col0, scale)
            ops.rope_kv_append(qkv_buf, a["q"], c.keys, c.values, lengths, B, L, at.n_heads,
                               at.n_kv_heads, at.head_dim, args.rope_theta, self.rope_scale,
                               at.q_norm.weight if at.q_norm else None, at.k_norm.weight if at.k_norm else None,
                               args.rms_norm_eps)
            ops.attn(a["q"], c.keys, c.values, lengths, a["att"], B, L, at.n_heads, at.n_kv_heads, at.head_dim,
                     at.scale, workspace=a["ws_attn"], out_perm=pm)
            if fx is not None:
                # row-parallel partial sums accumulate into the peer-mapped buffer; ONE kernel then does the
                # all-reduce over NVLink + residual add + post-attention RMSNorm (and re-zeroes the other buffer)
                at.o_proj(a["att"], EPI_ADD_RESIDUAL, out=fx.partial(0, M), workspace=a["ws"], x_perm=pm)
                n2 = layer.post_attention_layernorm
                if rows_mode:
                    fx.exchange_rows(0, a["h"], n2.weight, n2.eps, out_perm=pm)
                else:
                    fx.exchange(0, a["h"], n2.weight, a["x"], n2.eps, out_perm=pm)
            else:
                if tp and self.tp_rank != 0:
                    a["h"].zero_()
                at.o_proj(a["att"], EPI_ADD_RESIDUAL, out=a["h"], workspace=a["ws"], x_perm=pm)
                if tp:
                    self._allreduce(a["h"])
                ops.rmsnorm(a["h"], layer.post_attention_layernorm.weight, layer.post_attention_layernorm.eps, out=a["x"],
                            out_perm=pm)
            if a["gu32"] is not None:      # small (sharded) gate|up: fp32 accumulator, silu_mul rounds and re-zeroes it
                mlp.gate_up_proj(a["x"], EPI_ADD_RESIDUAL, out=a["gu32"], workspace=a["ws"], x_perm=pm)
                ops.silu_mul(a["gu32"], a["act"], out_perm=pm)
            else:
                mlp.gate_up_proj(a["x"], EPI_SILU_MUL, out=a["act"], workspace=a["ws"], x_perm=pm, y_perm=pm)
            if fx is not None:
                mlp.down_proj(a["act"], EPI_ADD_RESIDUAL, out=fx.partial(1, M), workspace=a["ws"], x_perm=pm)
                nxt = layers[li + 1].input_layernorm if li + 1 < len(layers) else None
                if rows_mode:
                    # the residual stream is row-sharded: the last exchange applies the FINAL norm so that x (the only
                    # tensor every rank holds completely) feeds the head directly
                    nn_ = nxt if nxt is not None else self.model.norm
                    fx.exchange_rows(1, a["h"], nn_.weight, nn_.eps, out_perm=pm)
                    a["x_final"] = nxt is None
                else:
                    fx.exchange(1, a["h"], nxt.weight if nxt else None, a["x"] if nxt else None,
                                nxt.eps if nxt else 0.0, out_perm=pm)
            else:
                if tp and self.tp_rank != 0:
                    a["h"].zero_()
                mlp.down_proj(a["act"], EPI_ADD_RESIDUAL, out=a["h"], workspace=a["ws"], x_perm=pm)
                if tp:
                    self._allreduce(a["h"])

    def _head(self, h_rows: Optional[torch.Tensor], a: dict, logits_out: Optional[torch.Tensor] = None,
              x_rows: Optional[torch.Tensor] = None):
        """final RMSNorm + lm_head / tied embed_tokens.as_linear (llama.py:231,249-252) -> float32 logits.
        ``x_rows``: rows already final-normalised (row-sharded TP exchange) -- the norm is skipped."""
        n = self.model.norm
        if x_rows is not None:
            x = x_rows
        else:
            rows = h_rows.shape[0]
            x = ops.rmsnorm(h_rows, n.weight, n.eps, out=a["xl"][:rows], out_perm=self._perm8)
        head = self.lm_head if self.lm_head is not None else self.model.embed_tokens
        return QuantizedLinear.__call__(head, x, EPI_STORE_F32, out=logits_out, workspace=a["ws"], x_perm=self._perm8)

    def __call__(self, inputs, cache: Optional[List[BatchedKVCache]] = None, last_only: bool = False):
        """inputs int [B, L] -> logits float32 [B, L, V] (or [B, 1, V] with last_only, the only rows
        generate_step reads, utils.py:404 / SURVEY B10).  Under TP the vocab axis is this rank's shard."""
        with self._lock:
            return self._forward_locked(inputs, cache, last_only)

    def _forward_locked(self, inputs, cache, last_only):
        if not isinstance(inputs, torch.Tensor):
            inputs = torch.as_tensor(inputs)
        inputs = inputs.to(self.device)
        B, L = inputs.shape
        ids = inputs.to(torch.int32).contiguous()
        own_cache = cache is None
        if own_cache:
            cache = self.make_cache(B, step=max(L, 1))
        prepare_cache(cache, B, L, self.dtype, self.device)
        Lc = max(1, min(L, MAX_TOKENS_PER_FORWARD // B))
        outs = []
        for l0 in range(0, L, Lc):
            l1 = min(L, l0 + Lc)
            n = l1 - l0
            a = self._arena(B * n, B)
            chunk = ids[:, l0:l1].contiguous().view(-1)
            self._layers_forward(chunk, B, n, cache, a)
            advance_cache(cache, n)
            src = a["x"] if a.get("x_final") else a["h"]      # row-sharded TP: x already holds final_norm(h) on every rank
            if last_only:
                if l1 == L:
                    last = src if n == 1 else src.view(B, n, -1)[:, -1, :].contiguous()
                    return (self._head(None, a, x_rows=last) if a.get("x_final") else self._head(last, a)).view(B, 1, -1)
            else:
                outs.append((self._head(None, a, x_rows=src) if a.get("x_final") else self._head(src, a)).view(B, n, -1))
        return outs[0] if len(outs) == 1 else torch.cat(outs, 1)


# ---------------------------------------------------------------------------------------------------
def bind_cache_group(caches: List[BatchedKVCache], device, step: Optional[int] = None):
    """Make all layers of one generation share a single device lengths tensor."""
    if not caches:
        return
    B = caches[0].batch_size
    shared = torch.zeros((B,), dtype=torch.int32, device=device)
    for c in caches:
        if step is not None:
            c.step = step
        c.share_lengths(shared)
        c._sync_lengths()


def prepare_cache(caches: List[BatchedKVCache], B: int, L: int, dtype, device) -> bool:
    """Ensure every layer can take L more tokens per row.  Returns True if any buffer moved."""
    moved = False
    for c in caches:
        assert c.batch_size == B, "KV cache batch size mismatch"
        need = (max(c._offsets) if c._offsets else 0) + L
        moved |= c.ensure_capacity(need, dtype, device)
        c.lengths(device)
    return moved


def advance_cache(caches: List[BatchedKVCache], L: int):
    """lengths += L once per distinct device tensor; host mirrors on every layer."""
    seen = set()
    for c in caches:
        t = c._lengths
        if t.data_ptr() not in seen:
            seen.add(t.data_ptr())
            ops.advance_lengths(t, L)
        c.advance_host(L)
