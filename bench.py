#!/usr/bin/env python
"""bench.py -- decode tok/s of the batched decode step (BASELINE.json metric), one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]

N == 1: Llama-3-8B architecture, MLX 4-bit group-64, batch 64, 8192-token KV context (BASELINE.json
configs[2], the configuration the metric is quoted on), random-init weights + synthetic KV (no network).
N  > 1: the same workload tensor-parallel over N ranks (strong scaling), launched by torch.distributed.run.

A "step" = one decode step of the whole batch: embed -> 32 blocks -> head -> argmax, KV append included.
  value   device-timed tok/s (CUDA events around K graph replays, inputs resident in HBM)
  e2e     the same step driven with HOST token buffers: pinned H2D of the step's ids + D2H of its result
  roofline  dominant kernel (by bytes) timed with CUDA events in an eager instrumented pass
  cpu_baseline  the oracle (numpy port of the reference path) on a bounded sample, rank 0, N == 1 only
--impl reference times that CPU port alone (MLX itself is not installable here: see DESIGN.md).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decode_tokens_per_sec"
UNIT = "tok/s"

WORKLOADS = {
    # name: (config, batch, context tokens)
    "llama3-8b-4bit-b64-t8192": ("llama-3-8b", 64, 8192),
    "llama3-8b-4bit-b64-t256": ("llama-3-8b", 64, 256),
    "llama3-8b-4bit-b8-t8192": ("llama-3-8b", 8, 8192),
    "llama3-8b-4bit-b1-t8192": ("llama-3-8b", 1, 8192),
    # configs[2]'s B=256 point: 256 x 8192 tokens of 16-bit KV are 275 GB and do not fit one GPU; context capped at 4096 (137 GB)
    "llama3-8b-4bit-b256-t4096": ("llama-3-8b", 256, 4096),
    "mistral-7b-4bit-b8-t256": ("mistral-7b", 8, 256),
    "qwen3-14b-4bit-b64-t256": ("qwen3-14b", 64, 256),
    # configs[4]: Llama-3-70B-class 8-bit, batch 128, TP=8.  The quoted 32k context needs 171.8 GB of 16-bit KV per GPU next
    # to 9.2 GB of weights and does not fit 180 GB; 24 576 tokens (128.8 GB of KV per GPU) is the largest 8k multiple that does
    "llama3-70b-8bit-b128-t24576": ("llama-3-70b", 128, 24576),
    "llama3-70b-8bit-b128-t256": ("llama-3-70b", 128, 256),
    "small": ("small-llama", 4, 64),
}
HEADLINE = "llama3-8b-4bit-b64-t8192"
DEFAULT_SWEEP = ["llama3-8b-4bit-b64-t256", "llama3-8b-4bit-b8-t8192", "llama3-8b-4bit-b1-t8192"]


def ncu_traffic(workload, kernel, tp_world):
    """DRAM bytes per launch of the dominant kernel (dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full`
    capture) looked up in profiles/ncu_traffic.json, which names the capture file it was read from.  A capture is not
    something this run can redo (a number taken under a profiler is never a bench value), so the line carries the committed
    measurement together with its source, or null when that (workload, kernel, tp) was never captured."""
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        return None, None
    e = table.get(f"{workload}|{kernel}|tp{tp_world}")
    return (float(e["bytes_per_launch"]), e["source"]) if e else (None, None)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------ our arm
def build_model(cfg_name, device, tp_rank=0, tp_world=1, seed=1234):
    import torch
    from mlx_parallm_b200 import synth
    from mlx_parallm_b200.utils import _get_classes
    cfg = synth.get_config(cfg_name)
    w = synth.random_weights(cfg, seed=seed, dtype="float16", device=device)
    if tp_world > 1:
        from mlx_parallm_b200 import tp
        w = tp.shard_weights(w, cfg, tp_rank, tp_world)
    Model, Args = _get_classes(cfg)
    model = Model(Args.from_dict(cfg)).load_weights(w, device=device)
    del w
    torch.cuda.empty_cache()
    return cfg, model


def make_session(model, B, T, extra, seed=7):
    """KV cache holding T synthetic context tokens per row (capacity T + extra) and a decode session."""
    import torch
    from mlx_parallm_b200.utils import DecodeSession
    cache = model.make_cache(B, step=T + extra)
    g = torch.Generator(device=model.device)
    g.manual_seed(seed)
    for c in cache:
        c.ensure_capacity(T + extra, model.dtype, model.device)
        # synthetic context: decode-step cost does not depend on the values; unit-scale keys keep softmax sane
        c.keys.normal_(0.0, 0.3, generator=g)
        c.values.normal_(0.0, 0.3, generator=g)
        c.offsets_list = [T] * B
        c.offset = T
    sess = DecodeSession(model, cache, B, greedy=True)
    sess.tok.copy_(torch.randint(0, model.args.vocab_size, (B,), generator=g, device=model.device, dtype=torch.int32))
    return cache, sess


def time_steps(sess, steps, warmup, dist=None):
    import torch
    for _ in range(warmup):
        sess.run()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sess.run()
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms / steps


def time_e2e(sess, steps, warmup, dist=None):
    """Same step through host buffers: pinned H2D of the ids, graph replay, D2H of the sampled ids, every step."""
    import torch
    B = sess.B
    h_in = torch.zeros((B,), dtype=torch.int32).pin_memory()
    h_out = torch.zeros((B,), dtype=torch.int32).pin_memory()
    h_in.copy_(sess.tok.cpu())

    def one():
        sess.tok.copy_(h_in, non_blocking=True)
        sess.run()
        h_out.copy_(sess.tok, non_blocking=True)
        torch.cuda.current_stream().synchronize()      # the caller reads the ids on the host
        h_in.copy_(h_out)

    for _ in range(warmup):
        one()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        one()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms / steps, 4 * B, 4 * B


def kernel_profile(sess, steps=3):
    """Eager (no graph) instrumented pass: CUDA events around every b200p launch -> mean ms per kernel kind and
    launches per step."""
    import torch
    from mlx_parallm_b200 import ops
    saved = sess.use_graph
    sess.use_graph = False
    sess.run()
    torch.cuda.synchronize()
    ops.PROFILER = {}
    n0 = ops.LAUNCHES
    for _ in range(steps):
        sess.run()
    torch.cuda.synchronize()
    prof, ops.PROFILER = ops.PROFILER, None
    launches = (ops.LAUNCHES - n0) // steps
    sess.use_graph = saved
    out = {}
    for k, evs in prof.items():
        ts = [s.elapsed_time(e) for s, e in evs]
        out[k] = {"calls_per_step": len(ts) // steps, "mean_ms": sum(ts) / len(ts), "total_ms_per_step": sum(ts) / steps}
    return out, launches


def run_workload(name, model, cfg, steps, warmup, tp_world, dist, with_profile, hbm_peak):
    import torch
    from mlx_parallm_b200 import synth
    _, B, T = WORKLOADS[name]
    extra = 2 * (warmup + steps) + 64
    cache, sess = make_session(model, B, T, extra)
    ms = time_steps(sess, steps, warmup, dist)
    t_mid = T + warmup + steps / 2.0
    w_bytes = synth.weight_bytes_per_step(cfg)
    kv_tok = synth.kv_bytes_per_token(cfg)
    step_bytes = w_bytes + B * t_mid * kv_tok
    res = {"workload": name, "batch": B, "context": T, "ms_per_step": ms, "tok_s": B / ms * 1e3,
           "bytes_per_step": step_bytes, "step_gbs_per_gpu": step_bytes / tp_world / ms / 1e6,
           "step_frac_of_hbm_peak": step_bytes / tp_world / ms / 1e6 / hbm_peak}
    if with_profile:
        e2e_ms, h2d, d2h = time_e2e(sess, steps, warmup, dist)
        res["e2e"] = {"ms_per_step": e2e_ms, "tok_s": B / e2e_ms * 1e3, "h2d": h2d, "d2h": d2h}
        prof, launches = kernel_profile(sess)
        res["kernels"] = prof
        res["launches_per_step"] = launches
        # dominant kernel by algorithmic bytes
        t_now = cache[0].offset
        attn_bytes = B * t_now * kv_tok / cfg["num_hidden_layers"] / tp_world        # per launch (one layer)
        kinds = {"b200p_attn": attn_bytes}
        inter, h = cfg["intermediate_size"], cfg["hidden_size"]
        bits, gs = cfg["quantization"]["bits"], cfg["quantization"]["group_size"]
        for k in prof:
            if k.startswith("b200p_qmm[N="):
                n_, k_ = (int(v.split("=")[1]) for v in k[len("b200p_qmm["):-1].split(","))
                kinds[k] = n_ * k_ * bits / 8 + 2 * n_ * (k_ // gs) * 2
        per_step = {k: kinds[k] * prof[k]["calls_per_step"] for k in kinds if k in prof}
        dom = max(per_step, key=per_step.get)
        ach = kinds[dom] / (prof[dom]["mean_ms"] * 1e-3) / 1e9
        # measured DRAM traffic of one launch of the dominant kernel (ncu --set full, dram__bytes_read.sum +
        # dram__bytes_write.sum): only known for the configuration that was captured
        traffic, traffic_src = ncu_traffic(name, dom, tp_world)
        exch = sum(v["total_ms_per_step"] for k, v in prof.items() if "allreduce" in k or "peer" in k or "exchange" in k)
        res["exchange_share_of_step_time"] = exch / sum(v["total_ms_per_step"] for v in prof.values())
        res["roofline"] = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                           "frac": ach / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                           "algorithmic_bytes_per_launch": kinds[dom],
                           "share_of_step_time": prof[dom]["total_ms_per_step"] / sum(v["total_ms_per_step"] for v in prof.values())}
    del sess, cache
    torch.cuda.empty_cache()
    return res


def time_prefill(model, cfg, B, Lp, reps=3, dist=None):
    """prefill tok/s = B * Lp / t_first_token (utils.py:533-537): the eager chunked prefill of a [B, Lp] prompt batch through
    the public Model.__call__ (last-position logits) + greedy sample, device-timed, max over ranks."""
    import torch
    g = torch.Generator(device=model.device)
    g.manual_seed(11)
    ids = torch.randint(0, cfg["vocab_size"], (B, Lp), generator=g, device=model.device)
    cache = model.make_cache(B, step=Lp + 8)
    times = []
    for r in range(reps + 1):
        for c in cache:
            c.reset(B)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        logits = model(ids, cache=cache, last_only=True)[:, -1, :]
        tok = torch.empty((B,), dtype=torch.int32, device=model.device)
        model.greedy_tokens(logits.contiguous(), tok)
        e1.record()
        torch.cuda.synchronize()
        if r > 0:
            times.append(e0.elapsed_time(e1))
    ms = min(times)
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    del cache
    torch.cuda.empty_cache()
    return {"batch": B, "prompt_len": Lp, "ms": ms, "tok_s": B * Lp / ms * 1e3}


def time_request(model, B=8, Lp=32, n_new=16, reps=2):
    """Latency of ONE short request through the public API (utils.generate_step: prefill, decode-graph capture or reuse, a host
    read of every token like the server's), wall clock: time to the first token and to the last of n_new.  The first
    repetition pays the capture of the decode graph for this (batch, cache group); later ones show the steady state."""
    import torch
    from mlx_parallm_b200.utils import generate_step
    g = torch.Generator(device="cpu").manual_seed(3)
    prompt = torch.randint(0, model.args.vocab_size, (B, Lp), generator=g)
    out = []
    for rep in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ttft = None
        for i, (tok, _) in zip(range(n_new), generate_step(prompt, model, temp=0.0)):
            tok.tolist()
            if ttft is None:
                ttft = time.perf_counter() - t0
        total = time.perf_counter() - t0
        out.append({"ttft_ms": ttft * 1e3, "total_ms": total * 1e3})
    return {"batch": B, "prompt_len": Lp, "new_tokens": n_new, "timing": "host wall clock, every token read on the host",
            "first_request": out[0], "later_request": out[-1]}


def tp_selfcheck(cfg_name, cfg, model, rank, world, local_rank, dist, B=8, L=32, steps=8):
    """Correctness of the sharded model carried by the multi-GPU line itself (outside the timed region): rank 0 also builds
    the UNSHARDED model from the same seed, every rank runs the same prefill + ``steps`` greedy steps through the sharded one,
    and rank 0 compares (i) its vocabulary shard of the prefill logits with the unsharded logits, (ii) the greedy ids.  A
    sequence may leave the unsharded transcript only where the unsharded logits put the sharded model's token within the
    1e-2 relative tolerance of the maximum (the exchange sums partial results in another order)."""
    import numpy as np
    import torch
    from mlx_parallm_b200.utils import generate_step
    g = torch.Generator(device="cpu").manual_seed(21)
    ids = torch.randint(0, cfg["vocab_size"], (B, L), generator=g)
    cache = model.make_cache(B, step=L + steps + 8)
    shard = model(ids, cache=cache, last_only=True)[:, -1].float().cpu().numpy()          # [B, V / world] on this rank
    del cache
    toks = np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(steps), generate_step(ids, model))], 1)
    out = None
    if rank == 0:
        _, ref = build_model(cfg_name, f"cuda:{local_rank}", 0, 1)
        rc = ref.make_cache(B, step=L + steps + 8)
        full = ref(ids, cache=rc, last_only=True)[:, -1].float().cpu().numpy()
        del rc
        vs = shard.shape[1]
        err = float(np.abs(shard - full[:, :vs]).max() / np.abs(full).max())
        ref_toks, ref_logits = [], []
        rc = ref.make_cache(B, step=L + steps + 8)
        lg = ref(ids, cache=rc, last_only=True)[:, -1]
        for _ in range(steps):
            y = lg.argmax(-1)
            ref_toks.append(y.cpu().numpy())
            ref_logits.append(lg.float().cpu().numpy())
            lg = ref(y[:, None], cache=rc, last_only=True)[:, -1]
        ref_toks = np.stack(ref_toks, 1)
        identical, near, real = 0, 0, 0
        for b in range(B):
            d = np.nonzero(toks[b] != ref_toks[b])[0]
            if d.size == 0:
                identical += 1
                continue
            s_ = int(d[0])
            l_ = ref_logits[s_][b]
            if float(l_.max() - l_[toks[b, s_]]) <= 1e-2 * float(np.abs(ref_logits[s_]).max()):
                near += 1
            else:
                real += 1
        out = {"prefill_logits_rel_err_vs_unsharded": err, "sequences": B, "greedy_steps": steps, "identical": identical,
               "left_at_a_near_tie": near, "real_divergences": real, "ok": bool(err < 1e-2 and real == 0)}
        del ref, rc
        torch.cuda.empty_cache()
    dist.barrier()
    return out


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_reference(cfg_name, B, T, steps=3, warmup=1, budget_layers=2, seed=1234):
    """The oracle (numpy restatement of the reference's decode step) timed on the host cores on a bounded sample.
    ONE SAMPLE STEP = ``budget_layers`` full-width transformer blocks + the head at the full batch and context; ``steps``
    sample steps are timed one by one after ``warmup`` untimed ones.  The full-model step is extrapolated from the fastest
    sample as t_layer * L + t_head (the reference's step is L identical blocks + head; the head alone is timed separately)."""
    import numpy as np
    import torch
    from mlx_parallm_b200 import synth
    from oracle.model import OracleConfig, OracleModel
    # use every host core: torchrun exports OMP_NUM_THREADS=1 to its workers, which would silently make this a 1-thread run
    ncores = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=ncores)
    except Exception:
        pass
    torch.set_num_threads(ncores)
    cfg = synth.get_config(cfg_name)
    L = cfg["num_hidden_layers"]
    sub = dict(cfg, num_hidden_layers=budget_layers)
    w = synth.random_weights(sub, seed=seed, dtype="float16", device="cpu")
    om = OracleModel(OracleConfig.from_hf(sub, "float16"), synth.to_numpy_weights(w), numerics="reference")
    rng = np.random.default_rng(7)
    cache = om.make_cache(B)
    for c in cache:
        c.step = T + 8 + steps + warmup
        c._ensure(T + 8 + steps + warmup)
        c.keys[:, :, :T] = rng.standard_normal((1, c.n_kv_heads, T, c.head_dim), dtype=np.float32) * 0.3
        c.values[:, :, :T] = rng.standard_normal((1, c.n_kv_heads, T, c.head_dim), dtype=np.float32) * 0.3
        c.offsets_list = [T] * B
    ids = rng.integers(0, cfg["vocab_size"], (B, 1))
    for _ in range(max(1, warmup)):                            # the first pass also builds the dequantised weight cache
        om.forward(ids, cache, last_only=True)
    samples = []
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        om.forward(ids, cache, last_only=True)
        samples.append(time.perf_counter() - t0)
    # head alone (a 0-layer model sharing the weights and the dequantised cache)
    om0 = OracleModel(OracleConfig.from_hf(dict(cfg, num_hidden_layers=0), "float16"), om.w, numerics="reference")
    om0._deq_cache = om._deq_cache
    heads = []
    for _ in range(3):
        t0 = time.perf_counter()
        om0.forward(ids, [], last_only=True)
        heads.append(time.perf_counter() - t0)
    t_all, t_head = min(samples), min(heads)
    t_layer = max(t_all - t_head, 1e-9) / budget_layers
    t_step = t_layer * L + t_head
    return {"value": B / t_step, "unit": UNIT, "cores": ncores, "kind": "port",
            "sample": f"one sample step = {budget_layers} of {L} blocks + head at B={B}, T={T}, numpy/BLAS threads={ncores}; "
                      f"{len(samples)} timed sample steps (min {t_all * 1e3:.0f} ms, max {max(samples) * 1e3:.0f} ms), "
                      f"t_layer={t_layer * 1e3:.1f} ms, t_head={t_head * 1e3:.1f} ms, full step extrapolated as t_layer*{L}+t_head",
            "sample_steps_timed": len(samples), "ms_per_sample_step": sum(samples) / len(samples) * 1e3,
            "ms_per_sample_step_min": t_all * 1e3, "ms_per_sample_step_max": max(samples) * 1e3,
            "ms_per_step_full_model": t_step * 1e3}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=HEADLINE)
    ap.add_argument("--sweep", default=",".join(DEFAULT_SWEEP), help="comma list of extra workloads ('' = none)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prefill", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE line (the JSON); anything libraries print (NCCL's version banner, warnings) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line: dict):
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    cfg_name, B, T = WORKLOADS[args.workload]

    if args.impl == "reference":
        if rank != 0:
            return
        for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):     # before numpy / BLAS load
            os.environ[var] = str(os.cpu_count() or 1)
        # K timed sample steps after W warm-up ones, as asked (capped so that the run stays within a few minutes: a sample
        # step of the headline workload takes ~1-2 s on 16 cores); `steps` / `ms_per_step` describe exactly what was timed,
        # `value` is the full-model extrapolation spelled out in cpu_baseline.sample
        k = max(1, min(args.steps, 60))
        cb = cpu_reference(cfg_name, B, T, steps=k, warmup=max(1, min(args.warmup, 3)))
        line = {"metric": METRIC, "value": cb["value"], "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
                "steps": cb["sample_steps_timed"], "warmup": max(1, min(args.warmup, 3)), "ms_per_step": cb["ms_per_sample_step"],
                "ms_per_step_full_model": cb["ms_per_step_full_model"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic (random-init MLX-format weights, synthetic KV context)",
                "config": {"workload": args.workload, "batch": B, "context": T, "parallelism": "host cores (tp1)",
                           "tp_exchange": None,
                           "l2_policy": "inputs larger than every cache (dequantised fp32 weights + fp32 KV are re-streamed every step)"},
                "note": "the reference's arithmetic lives in mlx==0.25.2 (not installable offline): this is the numpy port of the same step",
                "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return

    import torch
    from mlx_parallm_b200 import _lib, ops
    torch.cuda.set_device(local_rank)
    _lib.require_device()
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod
    hbm_peak, peak_src = peaks()
    cfg, model = build_model(cfg_name, f"cuda:{local_rank}", rank, world)
    if world > 1:
        from mlx_parallm_b200 import tp
        tp.attach(model, dist, rank, world)
    selfcheck = tp_selfcheck(cfg_name, cfg, model, rank, world, local_rank, dist) if world > 1 else None
    sampler = ClockSampler(local_rank)
    sampler.start()
    main_res = run_workload(args.workload, model, cfg, args.steps, args.warmup, world, dist, True, hbm_peak)
    clocks = sampler.stop()
    sweep = []
    for name in [s for s in args.sweep.split(",") if s and s != args.workload]:
        cn = WORKLOADS[name][0]
        if cn != cfg_name:
            del model
            torch.cuda.empty_cache()
            cfg, model = build_model(cn, f"cuda:{local_rank}", rank, world)
            if world > 1:
                tp.attach(model, dist, rank, world)
            cfg_name = cn
        try:
            r = run_workload(name, model, cfg, max(8, args.steps // 2), args.warmup, world, dist, False, hbm_peak)
            sweep.append({k: r[k] for k in ("workload", "ms_per_step", "tok_s", "step_frac_of_hbm_peak")})
        except Exception as ex:  # a sweep entry must not sink the headline line
            sweep.append({"workload": name, "error": str(ex)[:200]})
    prefill = None
    if not args.no_prefill:
        # configs[2] says "8k-token KV context, prefill + decode": short prompts (GEMM-bound) and 2k / 8k prompts (where the
        # causal attention grows quadratically), one sequence and eight
        prefill = []
        for pb, pl in ((8, 256), (1, 2048), (8, 2048), (1, 8192), (8, 8192)):
            try:
                prefill.append(time_prefill(model, cfg, pb, pl, reps=2, dist=dist))
            except Exception as ex:
                prefill.append({"batch": pb, "prompt_len": pl, "error": str(ex)[:200]})
    request = None
    if world == 1 and not args.no_prefill:
        try:
            request = time_request(model)
        except Exception as ex:
            request = {"error": str(ex)[:200]}
    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cb = cpu_reference(WORKLOADS[args.workload][0], B, T, steps=3, warmup=1)
        except Exception as ex:
            cb = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"[:200]}
    if rank == 0:
        rf = main_res["roofline"]
        line = {"metric": METRIC, "value": main_res["tok_s"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": main_res["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f16 storage / f32 accumulate (4-bit g64 weights)",
                "data": "synthetic (random-init MLX-format weights, synthetic KV context)",
                "config": {"workload": args.workload, "batch": B, "context": T, "parallelism": f"tp{world}",
                           "tp_exchange": (getattr(getattr(model, "_tp_fused", None), "mode", "nccl") if world > 1 else None),
                           "l2_policy": "inputs larger than L2 (weights 4.2 GB + KV >> 126 MB are re-streamed every step)"},
                "e2e": {"value": main_res["e2e"]["tok_s"], "unit": UNIT, "h2d_bytes_per_step": main_res["e2e"]["h2d"],
                        "d2h_bytes_per_step": main_res["e2e"]["d2h"], "ms_per_step": main_res["e2e"]["ms_per_step"]},
                "gpu_launches": main_res["launches_per_step"] * args.steps,
                "roofline": {k: rf[k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")},
                "roofline_detail": {"kernel": rf["kernel"], "peak_source": peak_src,
                                    "algorithmic_bytes_per_launch": rf["algorithmic_bytes_per_launch"],
                                    "share_of_step_time": rf["share_of_step_time"],
                                    "traffic_source": rf["traffic_source"],
                                    "exchange_share_of_step_time": main_res["exchange_share_of_step_time"],
                                    "step_bytes": main_res["bytes_per_step"],
                                    "step_frac_of_hbm_peak": main_res["step_frac_of_hbm_peak"],
                                    "step_frac_of_8TBs_nominal": main_res["step_gbs_per_gpu"] / 8000.0},
                "kernels": {k: {"calls": v["calls_per_step"], "mean_us": round(v["mean_ms"] * 1e3, 2)} for k, v in main_res["kernels"].items()},
                "clocks": clocks, "sweep": sweep, "prefill": prefill, "request_latency": request}
        if selfcheck is not None:
            line["tp_selfcheck"] = selfcheck
        if cb is not None:
            line["cpu_baseline"] = cb
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
