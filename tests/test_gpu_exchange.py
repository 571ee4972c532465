"""The peer-memory exchange kernels (csrc/allreduce.cu) on ONE device: W simulated ranks, each with its own buffers,
flag array and stream, launched concurrently (all CTAs are co-resident, so the cross-"rank" handshake completes).
The real multi-GPU run of the same kernels is scripts/tp_check.py under torchrun (profiles/*tp*)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TD = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def _perm8(n):
    return (n & ~7) | (((n & 3) << 1) | ((n >> 2) & 1))


def _rmsnorm_ref(h, w, eps, dt):
    """mx.fast.rms_norm as our kernels round it: fp32 x * rsqrt(mean x^2 + eps) * w -> storage type."""
    inv = 1.0 / torch.sqrt((h.double() ** 2).mean(-1, keepdim=True).float() + eps)
    return (h * inv * w.float()[None, :]).to(dt)


@pytest.mark.parametrize("W,M,H,dtype,perm", [(2, 64, 4096, "float16", True), (4, 64, 5120, "bfloat16", False),
                                              (2, 5, 256, "float32", False), (4, 1500, 512, "float16", True),
                                              (8, 8, 1024, "float16", False)])
def test_allreduce_residual_rmsnorm_simulated_ranks(W, M, H, dtype, perm):
    from mlx_parallm_b200 import _lib
    L = _lib.lib()
    _lib.require_device()
    dt = TD[dtype]
    # all W "ranks" share this one device: keep W x grid CTAs co-resident or the handshake cannot complete
    _lib.check(L.b200p_allreduce_set_max_ctas(128 // W if M > 64 else 0))
    g = torch.Generator(device="cuda").manual_seed(5)
    flag_words = int(L.b200p_allreduce_flag_bytes()) // 4
    flags = [torch.zeros((flag_words,), dtype=torch.int32, device="cuda") for _ in range(W)]
    bufs = [[torch.zeros((M, H), dtype=torch.float32, device="cuda") for _ in range(W)] for _ in range(2)]   # A / B per rank
    nw = (1.0 + 0.1 * torch.randn((H,), generator=g, device="cuda")).to(dt)
    h0 = torch.randn((M, H), generator=g, device="cuda")
    hs = [h0.clone() for _ in range(W)]
    ys = [torch.zeros((M, H), dtype=dt, device="cuda") for _ in range(W)]
    streams = [torch.cuda.Stream() for _ in range(W)]
    arr = C.c_void_p * W
    ftab = arr(*[f.data_ptr() for f in flags])
    h_ref = h0.clone()
    for it in range(4):                       # A, B, A, B: epochs advance, the idle buffer is re-zeroed each time
        which = it & 1
        parts = [torch.randn((M, H), generator=g, device="cuda") * (0.5 + r) for r in range(W)]
        for r in range(W):
            assert float(bufs[which][r].abs().max()) == 0.0, "the exchange must hand back a zeroed buffer"
            bufs[which][r].copy_(parts[r])
            bufs[1 - which][r].fill_(7.0)     # pretend the other projection left partial sums behind
        torch.cuda.synchronize()
        ptab = arr(*[b.data_ptr() for b in bufs[which]])
        with_norm = it != 3
        for r in range(W):
            with torch.cuda.stream(streams[r]):
                rc = L.b200p_allreduce_residual_rmsnorm(ptab, ftab, W, r, hs[r].data_ptr(), nw.data_ptr() if with_norm else 0,
                                                        ys[r].data_ptr() if with_norm else 0, bufs[1 - which][r].data_ptr(), M, H, H,
                                                        1e-5, {"float32": 0, "float16": 1, "bfloat16": 2}[dtype],
                                                        _lib.OUT_PERM8 if (perm and with_norm) else 0, streams[r].cuda_stream)
                _lib.check(rc, "b200p_allreduce_residual_rmsnorm")
        torch.cuda.synchronize()
        for p in parts:                       # rank order, fp32: what every rank must hold bit for bit
            h_ref = h_ref + p
        for r in range(W):
            out = C.c_uint32(9)
            _lib.check(L.b200p_peer_status(flags[r].data_ptr(), C.byref(out), 0), "peer_status")
            assert out.value == 0, "handshake timed out"
            assert torch.equal(hs[r], h_ref), f"rank {r} residual stream differs (iteration {it})"
            assert float(bufs[1 - which][r].abs().max()) == 0.0
            if with_norm:
                y_ref = _rmsnorm_ref(h_ref, nw, 1e-5, dt)
                if perm:
                    idx = torch.tensor([_perm8(n) for n in range(H)], device="cuda")
                    y_perm = torch.empty_like(y_ref)
                    y_perm[:, idx] = y_ref
                    y_ref = y_perm
                err = (ys[r].float() - y_ref.float()).abs().max().item()
                tol = {"float32": 1e-5, "float16": 2e-3, "bfloat16": 1.6e-2}[dtype] * max(1.0, y_ref.float().abs().max().item())
                assert err <= tol, (r, it, err, tol)
                assert torch.equal(ys[r], ys[0]), "ranks must produce identical normalised rows"
            bufs[which][r].zero_()            # stands in for the NEXT exchange's re-zeroing of this buffer
    _lib.check(L.b200p_allreduce_set_max_ctas(0))


def test_peer_allgather_simulated_ranks():
    from mlx_parallm_b200 import _lib
    L = _lib.lib()
    W, B = 4, 37
    flag_words = int(L.b200p_allreduce_flag_bytes()) // 4
    flags = [torch.zeros((flag_words,), dtype=torch.int32, device="cuda") for _ in range(W)]
    src = [torch.randn((B, 4), device="cuda") + r for r in range(W)]
    dst = [torch.zeros((W, B, 4), device="cuda") for _ in range(W)]
    streams = [torch.cuda.Stream() for _ in range(W)]
    arr = C.c_void_p * W
    ftab, stab = arr(*[f.data_ptr() for f in flags]), arr(*[s.data_ptr() for s in src])
    torch.cuda.synchronize()
    for rep in range(2):
        for r in range(W):
            with torch.cuda.stream(streams[r]):
                _lib.check(L.b200p_peer_allgather(stab, ftab, W, r, dst[r].data_ptr(), B * 16, streams[r].cuda_stream), "allgather")
        torch.cuda.synchronize()
        for r in range(W):
            assert torch.equal(dst[r], torch.stack(src, 0))


def test_exchange_argument_errors():
    from mlx_parallm_b200 import _lib
    L = _lib.lib()
    arr = C.c_void_p * 8
    t = torch.zeros((16,), device="cuda")
    tab = arr(*([t.data_ptr()] * 8))
    assert L.b200p_allreduce_residual_rmsnorm(tab, tab, 3, 0, t.data_ptr(), 0, 0, 0, 1, 4, 4, 1e-5, 0, 0, 0) == -1      # world 3
    assert L.b200p_allreduce_residual_rmsnorm(tab, tab, 2, 2, t.data_ptr(), 0, 0, 0, 1, 4, 4, 1e-5, 0, 0, 0) == -1      # rank >= world
    assert L.b200p_allreduce_residual_rmsnorm(tab, tab, 2, 0, t.data_ptr(), t.data_ptr(), 0, 0, 1, 4, 4, 1e-5, 0, 0, 0) == -1   # w without y
    assert L.b200p_allreduce_residual_rmsnorm(tab, tab, 2, 0, t.data_ptr(), 0, 0, 0, 0, 4, 4, 1e-5, 0, 0, 0) == 0       # M == 0: no launch


@pytest.mark.parametrize("W,M,H,dtype,perm", [(2, 64, 4096, "float16", True), (4, 64, 5120, "bfloat16", False),
                                              (8, 64, 4096, "float16", True), (4, 3, 256, "float32", False),
                                              (2, 700, 512, "float16", True), (8, 1, 1024, "float16", False)])
def test_allreduce_rows_rmsnorm_simulated_ranks(W, M, H, dtype, perm):
    """Two-shot, row-sharded exchange: every rank's x holds rms_norm(h + sum partials) for ALL rows, each rank's h only
    its own rows; the idle partial buffer is re-zeroed; repeated calls advance the epochs."""
    from mlx_parallm_b200 import _lib
    L = _lib.lib()
    dt = TD[dtype]
    rpr = (M + W - 1) // W
    _lib.check(L.b200p_allreduce_set_max_ctas(max(1, 128 // W) if rpr > 128 // W else 0))
    g = torch.Generator(device="cuda").manual_seed(11)
    flag_words = int(L.b200p_allreduce_flag_bytes()) // 4
    flags = [torch.zeros((flag_words,), dtype=torch.int32, device="cuda") for _ in range(W)]
    bufs = [[torch.zeros((M, H), dtype=torch.float32, device="cuda") for _ in range(W)] for _ in range(2)]
    xs = [torch.zeros((M, H), dtype=dt, device="cuda") for _ in range(W)]
    nw = (1.0 + 0.1 * torch.randn((H,), generator=g, device="cuda")).to(dt)
    h0 = torch.randn((M, H), generator=g, device="cuda")
    hs = [h0.clone() for _ in range(W)]
    streams = [torch.cuda.Stream() for _ in range(W)]
    arr = C.c_void_p * W
    ftab, xtab = arr(*[f.data_ptr() for f in flags]), arr(*[x.data_ptr() for x in xs])
    h_ref = h0.clone()
    code = {"float32": 0, "float16": 1, "bfloat16": 2}[dtype]
    for it in range(3):
        which = it & 1
        parts = [torch.randn((M, H), generator=g, device="cuda") * (0.5 + r) for r in range(W)]
        for r in range(W):
            bufs[which][r].copy_(parts[r])
            bufs[1 - which][r].fill_(3.0)
            xs[r].zero_()
        torch.cuda.synchronize()
        ptab = arr(*[b.data_ptr() for b in bufs[which]])
        for r in range(W):
            with torch.cuda.stream(streams[r]):
                _lib.check(L.b200p_allreduce_rows_rmsnorm(ptab, xtab, ftab, W, r, hs[r].data_ptr(), nw.data_ptr(),
                                                          bufs[1 - which][r].data_ptr(), M, H, H, 1e-5, code,
                                                          _lib.OUT_PERM8 if perm else 0, streams[r].cuda_stream), "rows exchange")
        torch.cuda.synchronize()
        for p in parts:
            h_ref = h_ref + p
        y_ref = _rmsnorm_ref(h_ref, nw, 1e-5, dt)
        if perm:
            idx = torch.tensor([_perm8(n) for n in range(H)], device="cuda")
            y_perm = torch.empty_like(y_ref)
            y_perm[:, idx] = y_ref
            y_ref = y_perm
        tol = {"float32": 1e-5, "float16": 2e-3, "bfloat16": 1.6e-2}[dtype] * max(1.0, y_ref.float().abs().max().item())
        for r in range(W):
            out = C.c_uint32(9)
            _lib.check(L.b200p_peer_status(flags[r].data_ptr(), C.byref(out), 0), "peer_status")
            assert out.value == 0, "handshake timed out"
            lo, hi = min(M, r * rpr), min(M, (r + 1) * rpr)
            assert torch.equal(hs[r][lo:hi], h_ref[lo:hi]), f"rank {r}: owned rows of h differ (iteration {it})"
            assert float(bufs[1 - which][r].abs().max()) == 0.0
            assert torch.equal(xs[r], xs[0]), "every rank must hold the same x"
            assert (xs[r].float() - y_ref.float()).abs().max().item() <= tol
            hs[r].copy_(h_ref)                # the model never reads stale rows; keep the reference simple
    _lib.check(L.b200p_allreduce_set_max_ctas(0))
