"""Tensor-parallel parity on real GPUs (run under torchrun, one rank per GPU):

every rank builds its shard of a small synthetic model, the group decodes greedily for a few steps and rank 0 compares the
tokens and the last-position logits with the CPU oracle running the UNSHARDED model.  B200P_TP_FUSED=0 forces the NCCL
all-reduce, 1 (default) uses the peer-memory all-reduce + residual + RMSNorm kernel."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch
import torch.distributed as dist


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from helpers import prompts_for, rel_err
    from mlx_parallm_b200 import _lib, synth, tp
    from mlx_parallm_b200.utils import _get_classes, generate_step
    from oracle.model import OracleConfig, OracleModel, generate_step as oracle_step
    _lib.require_device()
    fails = 0
    cases = [("small-llama", dict(num_key_value_heads=world, num_attention_heads=2 * world, hidden_size=256 * world,
                                  intermediate_size=512 * world), "float16", 4, 24),
             ("tiny-qwen3", dict(num_key_value_heads=world, num_attention_heads=2 * world, hidden_size=128 * world,
                                 intermediate_size=64 * world * 3), "bfloat16", 3, 9),
             ("small-llama", dict(num_key_value_heads=world, num_attention_heads=2 * world, hidden_size=256 * world,
                                  intermediate_size=512 * world), "float16", 64, 200)]
    if os.environ.get("TP_CHECK_QUICK"):
        cases = cases[:2]
    for fused in ([True, False] if os.environ.get("B200P_TP_FUSED", "1") != "0" else [False]):
        for name, ov, dtype, B, Lp in cases:
            cfg = synth.get_config(name, **ov)
            w = synth.random_weights(cfg, seed=1234, dtype=dtype, device="cpu")
            Model, Args = _get_classes(cfg)
            gm = Model(Args.from_dict(cfg)).load_weights(tp.shard_weights(w, cfg, rank, world), device=f"cuda:{local}")
            tp.attach(gm, dist, rank, world, fused=fused)
            p = prompts_for(cfg, B, Lp)
            steps = 6
            got = [y[:, 0].cpu().numpy() for _, (y, _) in zip(range(steps), generate_step(torch.from_numpy(p), gm))]
            lg = gm(torch.from_numpy(p), cache=gm.make_cache(B), last_only=True)[:, -1, :]
            allg = [torch.empty_like(lg) for _ in range(world)]
            dist.all_gather(allg, lg.contiguous())
            lg = torch.cat(allg, 1).cpu().numpy()
            status = gm._tp_fused.status() if getattr(gm, "_tp_fused", None) is not None else 0
            if rank == 0:
                om = OracleModel(OracleConfig.from_hf(cfg, dtype), synth.to_numpy_weights(w), numerics="reference")
                exp = [y[:, 0] for _, (y, _) in zip(range(steps), oracle_step(p, om))]
                lo = om.forward(p, om.make_cache(B), last_only=True)[:, -1, :]
                err = rel_err(lg, lo)
                same = float(np.mean(np.all(np.stack(got, 1) == np.stack(exp, 1), axis=1)))
                # bf16 on a random tiny model has near-tied logits: a flipped row is rounding, not a TP error (the NCCL and
                # fused paths flip the same rows); the f16 cases must match the oracle's ids on every row
                ok = err < 1e-2 and same >= (0.99 if dtype == "float16" else 0.6) and status == 0
                fails += 0 if ok else 1
                print(f"[tp{world} fused={int(fused)}] {name} {dtype} B={B} Lp={Lp}: logits rel err {err:.2e}, "
                      f"rows with identical greedy ids {same:.3f}, peer status {status} -> {'ok' if ok else 'FAIL'}", flush=True)
            del gm
            torch.cuda.empty_cache()
            dist.barrier()
    t = torch.tensor([fails], device="cuda")
    dist.broadcast(t, 0)
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()
