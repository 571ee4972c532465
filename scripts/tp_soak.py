"""Value-checking soak of the peer-memory TP exchange on real GPUs (torchrun, one rank per GPU).

The exchange kernels (csrc/allreduce.cu) publish their arrival with a relaxed flag and read the peers' partial buffers with
system-scope relaxed loads, relying on the kernel boundary before them for the writes to be performed.
This does not argue that; it measures it: >= 100 000 exchanges with random partial sums, every one of which must deliver

  * the bit-identical result on every rank (one-shot: the whole residual stream h and the normalised y; rows: the
    normalised x every rank receives), compared through an all-gather after every graph replay, and
  * the right VALUE: a float64 reference of sum-over-ranks accumulated over the replay (a partial read before it was
    complete, or zeroed early, is off by O(1), not by rounding),

with the status word (handshake time-outs) still zero at the end.  Exchanges run back to back inside CUDA graphs of 256 (the
way a decode step issues them), each preceded on the same stream by the copy that plays the row-parallel GEMM's epilogue.

    torchrun --nproc-per-node 8 scripts/tp_soak.py [exchanges=102400] [M=64] [H=4096]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    total = int(sys.argv[1]) if len(sys.argv) > 1 else 102400
    M = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    H = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from mlx_parallm_b200 import _lib, tp
    _lib.require_device()
    NX, K = 256, 4                                   # exchanges per graph, distinct partial patterns per rank
    ok_all = True
    for mode in ("oneshot", "rows"):
        fx = tp.PeerExchange(dist, None, rank, world, f"cuda:{local}", max(M, 64), H, mode=mode)
        # every rank can regenerate every rank's patterns (for the reference); pats[r][j]
        pats = [[torch.randn((M, H), generator=torch.Generator(device="cuda").manual_seed(1000 * r + j), device="cuda") for j in range(K)]
                for r in range(world)]
        mine = pats[rank]
        w = (torch.rand((H,), generator=torch.Generator(device="cuda").manual_seed(5), device="cuda") + 0.5).to(torch.float16)
        h = torch.zeros((M, H), dtype=torch.float32, device="cuda")
        y = torch.zeros((M, H), dtype=torch.float16, device="cuda")
        rows_per = (M + world - 1) // world
        r0, r1 = min(M, rank * rows_per), min(M, (rank + 1) * rows_per)

        def body():
            h.zero_()
            for i in range(NX):
                fx.partial(i & 1, M).copy_(mine[i % K])            # stands for the row-parallel GEMM's epilogue
                if mode == "oneshot":
                    fx.exchange(i & 1, h, w, y, 1e-5, out_perm=False)
                else:
                    fx.exchange_rows(i & 1, h, w, 1e-5, out_perm=False)

        body()
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        # float64 reference of one replay: h = sum_i sum_r pats[r][i % K], y = rms_norm(h) * w
        s = torch.zeros((M, H), dtype=torch.float64, device="cuda")
        for i in range(NX):
            for r in range(world):
                s += pats[r][i % K].double()
        ref_y = (s * torch.rsqrt((s * s).mean(-1, keepdim=True) + 1e-5) * w.double())
        replays = (total // 2 + NX - 1) // NX
        bad_bits = bad_val = 0
        worst = 0.0
        for it in range(replays):
            g.replay()
            out = y if mode == "oneshot" else fx.x_view(M, torch.float16)
            # (1) bit-identical on every rank
            gathered = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(gathered, out.contiguous())
            if any(not torch.equal(gathered[0].view(torch.int16), t.view(torch.int16)) for t in gathered[1:]):
                bad_bits += 1
            if mode == "oneshot":
                hg = [torch.empty_like(h) for _ in range(world)]
                dist.all_gather(hg, h)
                if any(not torch.equal(hg[0].view(torch.int32), t.view(torch.int32)) for t in hg[1:]):
                    bad_bits += 1
                err_h = float((h.double() - s).abs().max() / s.abs().max())
            else:
                err_h = float((h[r0:r1].double() - s[r0:r1]).abs().max() / s.abs().max()) if r1 > r0 else 0.0
            # (2) the right value (f16 output rounding: 1e-3 relative; fp32 accumulation of 256 x W terms: 1e-5)
            err_y = float((out.double() - ref_y).abs().max() / ref_y.abs().max())
            worst = max(worst, err_h, err_y / 100)
            if err_h > 1e-5 or err_y > 2e-3:
                bad_val += 1
        st = fx.status()
        res = torch.tensor([bad_bits, bad_val, st], device="cuda", dtype=torch.int64)
        dist.all_reduce(res)
        ok = int(res.sum().item()) == 0
        ok_all &= ok
        if rank == 0:
            print(f"tp{world} {mode}: {replays * NX} exchanges of {M} x {H} in graphs of {NX}: ranks bit-identical "
                  f"{'yes' if int(res[0]) == 0 else 'NO (' + str(int(res[0])) + ' replays)'}, values "
                  f"{'ok' if int(res[1]) == 0 else 'WRONG (' + str(int(res[1])) + ' replays)'} "
                  f"(worst residual-stream error {worst:.1e}), status word {int(res[2])} -> {'PASS' if ok else 'FAIL'}", flush=True)
        del g
        fx.close()
        dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0 if ok_all else 1)      # skip NCCL teardown: communicator destruction after graph-captured work can hang


if __name__ == "__main__":
    main()
