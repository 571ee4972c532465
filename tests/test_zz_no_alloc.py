"""b200p_attn_decode_static: decode attention whose split-K scratch is a file-scope device array of the library (no
cudaMalloc anywhere in the call), checked through the C-ABI against b200p_attn with a caller-allocated workspace.

Collected last on purpose; passes alone in a fresh process and after the other tests:
it keeps no state besides the attention path switch, which it restores."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu

TOL = {torch.float16: 2e-3, torch.bfloat16: 1.2e-2}     # the decode-attention tolerances of tests/test_gpu_kernels.py
SCRATCH_BYTES = 16 << 20                                # B200P_ATTN_STATIC_SCRATCH_BYTES (include/b200p.h)


@pytest.fixture(scope="module")
def L():
    from mlx_parallm_b200 import _lib
    _lib.require_device()
    return _lib.lib()


def _case(B, Hq, Hkv, lengths, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    D, Tcap = 128, int(max(lengths)) + 4
    q = torch.randn((B, 1, Hq, D), generator=g, device="cuda").to(dtype)
    k = torch.randn((B, Hkv, Tcap, D), generator=g, device="cuda").to(dtype)
    v = torch.randn((B, Hkv, Tcap, D), generator=g, device="cuda").to(dtype)
    ln = torch.tensor(lengths, dtype=torch.int32, device="cuda")
    return q, k, v, ln, Tcap


def _both(L, B, Hq, Hkv, lengths, dtype, seed, flags=0):
    """-> (output of b200p_attn_decode_static, output of b200p_attn on the tensor decode path), same inputs."""
    from mlx_parallm_b200 import _lib
    q, k, v, ln, Tcap = _case(B, Hq, Hkv, lengths, dtype, seed)
    D, scale, code = 128, 128 ** -0.5, _lib.F16 if dtype == torch.float16 else _lib.BF16
    st = torch.cuda.current_stream().cuda_stream
    out_s = torch.full((B, 1, Hq * D), float("nan"), dtype=dtype, device="cuda")
    out_r = torch.full((B, 1, Hq * D), float("nan"), dtype=dtype, device="cuda")
    nbytes = int(L.b200p_attn_workspace_bytes(B, 1, Hq, D))
    ws = torch.empty((max(nbytes, 16),), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(L.b200p_attn_decode_static(q.data_ptr(), k.data_ptr(), v.data_ptr(), ln.data_ptr(), out_s.data_ptr(), B, Hq, Hkv,
                                          D, Tcap, scale, code, flags, st), "b200p_attn_decode_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    _lib.check(L.b200p_attn_set_path(2))
    try:
        _lib.check(L.b200p_attn(q.data_ptr(), k.data_ptr(), v.data_ptr(), ln.data_ptr(), out_r.data_ptr(), B, 1, Hq, Hkv, D, Tcap,
                                scale, code, flags, ws.data_ptr(), ws.numel(), st), "b200p_attn")
    finally:
        L.b200p_attn_set_path(0)
    torch.cuda.synchronize()
    return out_s.float().cpu().numpy(), out_r.float().cpu().numpy(), free0 - free1


def test_static_scratch_decode_attention_matches_b200p_attn_twice_in_a_row(L):
    # both calls split the sequence across CTAs, so the partials really travel through the static array; the second call has
    # another batch, head layout, dtype and much shorter rows: anything it picked up from the first call's records would show
    calls = [
        (2, 8, 1, [5000, 37], torch.float16, 11),
        (3, 32, 8, [700, 3, 0], torch.bfloat16, 12),
    ]
    for i, (B, Hq, Hkv, lengths, dtype, seed) in enumerate(calls):
        got, exp, grew = _both(L, B, Hq, Hkv, lengths, dtype, seed)
        assert np.isfinite(got).all() and np.isfinite(exp).all()
        assert rel_err(got, exp) < TOL[dtype], f"call {i}"
        assert np.array_equal(got, exp), f"call {i}: same kernels and split count, the results should be bit-identical"
        if i > 0:     # the first call of a process may load the module; after that a call must not take device memory
            assert grew <= 0, f"call {i} reduced free device memory by {grew} bytes"


def test_static_scratch_perm8_output_and_empty_batch(L):
    from mlx_parallm_b200 import _lib
    got, exp, _ = _both(L, 2, 8, 2, [1500, 9], torch.float16, 13, flags=_lib.OUT_PERM8)
    assert np.array_equal(got, exp)
    assert L.b200p_attn_decode_static(1, 1, 1, 1, 1, 0, 8, 2, 128, 16, 1.0, _lib.F16, 0, None) == 0     # B == 0: no launch


def test_static_scratch_rejects_sizes_beyond_the_array(L):
    from mlx_parallm_b200 import _lib
    per_row_head = int(L.b200p_attn_workspace_bytes(1, 1, 1, 128))
    max_bh = SCRATCH_BYTES // per_row_head
    assert max_bh == 992
    q, k, v, ln, Tcap = _case(1, 8, 2, [5], torch.float16, 14)
    args = lambda B, Hq: (q.data_ptr(), k.data_ptr(), v.data_ptr(), ln.data_ptr(), q.data_ptr(), B, Hq, Hq // 4, 128, Tcap, 1.0,
                          _lib.F16, 0, None)
    # validation happens before anything is launched: the pointers are never dereferenced for a rejected call
    assert L.b200p_attn_decode_static(*args(max_bh // 8 + 1, 8)) == -1
    msg = _lib.last_error()
    assert "static" in msg and str(SCRATCH_BYTES) in msg and "b200p_attn" in msg
    assert L.b200p_attn_decode_static(*args(1, 8)[:8], 64, *args(1, 8)[9:]) == -1           # head_dim 64: unsupported
    assert L.b200p_attn_decode_static(*args(1, 8)[:11], _lib.F32, 0, None) == -1            # float32 caches: unsupported
    assert L.b200p_attn_decode_static(*args(1, 8)[:12], 0x80, None) == -1                   # unknown flag
