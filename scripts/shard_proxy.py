"""One GPU's share of a tensor-parallel decode step, timed WITHOUT the exchange (1 GPU is enough):

builds rank 0's shard of a world-W model (weights generated tensor by tensor on the device and sharded as they come, so
the unsharded model never has to exist), fills a synthetic KV context and times the captured decode step.  What it
measures is the per-GPU compute + HBM part of BASELINE configs[4] (Llama-3-70B-class 8-bit, TP=8), which does not fit
a box budget as an 8-GPU run: step time under TP = this + 2 exchanges per layer (profiles/r1_run10_tp8_exchange.txt).
Partial sums are NOT reduced, so the tokens are meaningless; the cost of a step does not depend on the values.

    python scripts/shard_proxy.py --config llama-3-70b --world 8 --batch 128 --context 16384
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="llama-3-70b")
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--context", type=int, default=16384)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    import torch
    import bench
    from mlx_parallm_b200 import _lib, synth, tp
    from mlx_parallm_b200.utils import _get_classes
    _lib.require_device()
    dev = "cuda:0"
    cfg = synth.get_config(args.config)
    tp.check_divisible(cfg, args.world)
    shard = {}
    prefixes = ["model.embed_tokens", "lm_head", "model.norm"] + [f"model.layers.{i}." for i in range(cfg["num_hidden_layers"])]
    for p in prefixes:
        w = synth.random_weights(cfg, seed=1234, dtype="float16", device=dev, only=lambda n, p=p: n.startswith(p))
        shard.update(tp.shard_weights(w, cfg, 0, args.world))
        del w
    Model, Args = _get_classes(cfg)
    model = Model(Args.from_dict(cfg)).load_weights(shard, device=dev)
    del shard
    torch.cuda.empty_cache()
    B, T = args.batch, args.context
    extra = 2 * (args.warmup + args.steps) + 64
    cache, sess = bench.make_session(model, B, T, extra)
    ms = bench.time_steps(sess, args.steps, args.warmup)
    prof, launches = bench.kernel_profile(sess)
    hbm_peak, src = bench.peaks()
    w_bytes = synth.weight_bytes_per_step(cfg) / args.world
    kv_bytes = B * (T + args.warmup + args.steps / 2.0) * synth.kv_bytes_per_token(cfg) / args.world
    # the embedding table is replicated, not sharded, but only B rows of it are read: weight_bytes_per_step already leaves it out
    gbs = (w_bytes + kv_bytes) / ms / 1e6
    out = {"workload": f"{args.config} rank-0 shard of TP={args.world}, B={B}, T={T}, no exchange", "ms_per_step": ms,
           "tok_s_if_exchange_were_free": B / ms * 1e3, "bytes_per_gpu_step": w_bytes + kv_bytes, "weights_gb": w_bytes / 1e9,
           "kv_gb": kv_bytes / 1e9, "gbs": gbs, "frac_of_hbm_peak": gbs / hbm_peak, "peak": hbm_peak, "peak_source": src,
           "launches_per_step": launches,
           "kernels": {k: {"calls": v["calls_per_step"], "mean_us": round(v["mean_ms"] * 1e3, 2),
                           "ms_per_step": round(v["total_ms_per_step"], 3)} for k, v in prof.items()}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
