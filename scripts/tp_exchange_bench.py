"""Microbenchmark of the per-layer TP exchange on real GPUs (torchrun, one rank per GPU): CUDA-graph of 128 exchanges,
device-timed, max over ranks.  fused  = b200p_allreduce_residual_rmsnorm (peer memory, one kernel)
                               nccl   = memset-free equivalent: ncclAllReduce(fp32 [M,H]) + b200p_rmsnorm."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from mlx_parallm_b200 import _lib, ops, tp
    _lib.require_device()
    NX = 128
    for M, H in [(64, 4096), (8, 4096), (1, 4096), (128, 8192), (2048, 4096)]:
        fx = tp.PeerExchange(dist, None, rank, world, f"cuda:{local}", max(M, 64), H)
        h = torch.zeros((M, H), dtype=torch.float32, device="cuda")
        y = torch.zeros((M, H), dtype=torch.float16, device="cuda")
        w = torch.ones((H,), dtype=torch.float16, device="cuda")
        n = NX if M <= 128 else 16

        def fused_body():
            for i in range(n):
                fx.exchange(i & 1, h, w, y, 1e-5, out_perm=True)

        def rows_body():
            for i in range(n):
                fx.exchange_rows(i & 1, h, w, 1e-5, out_perm=True)

        def nccl_body():
            for i in range(n):
                dist.all_reduce(h)
                ops.rmsnorm(h, w, 1e-5, out=y, out_perm=True)

        res = {}
        for name, body in (("fused", fused_body), ("rows", rows_body), ("nccl", nccl_body)):
            body()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                body()
            res[name] = timed(g.replay) * 1e3 / n
            h.zero_()
        st = fx.status()
        if rank == 0:
            byt = world * M * H * 4
            print(f"tp{world} M={M} H={H}: one-shot {res['fused']:.2f} us/exchange ({byt / res['fused'] / 1e3:.0f} GB/s peer reads), "
                  f"rows (two-shot) {res['rows']:.2f} us, "
                  f"nccl all-reduce + rmsnorm {res['nccl']:.2f} us, status {st}", flush=True)
        del g
        fx.close()
        dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)          # skip NCCL teardown: communicator destruction after graph-captured collectives can hang


if __name__ == "__main__":
    main()
