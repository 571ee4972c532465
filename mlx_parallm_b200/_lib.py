"""ctypes binding of libb200parallm.so (the C-ABI declared in include/b200p.h).

There is no fallback: if the shared library is missing or the device is not sm_100 every call
raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C mlx_parallm_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200parallm.so")
# profiling builds only (scripts/tc_bisect.py): a library compiled with -DB200P_TC_DEBUG=<mask> under build/
LIB_PATH = os.environ.get("B200P_LIB", LIB_PATH)

F32, F16, BF16 = 0, 1, 2
DQ_F32_MULADD, DQ_FMA, DQ_T_MULADD = 0, 1, 2
EPI_STORE, EPI_STORE_F32, EPI_ADD_RESIDUAL, EPI_SILU_MUL = 0, 1, 2, 3
X_PERM8, Y_PERM8, OUT_PERM8 = 0x100, 0x200, 0x1      # activation-layout flags (include/b200p.h)
W_SMALLSCALE = 0x400                                 # every |scale| < 0.0625: subnormal-operand dequant (f16 only)

_vp, _i64, _i32, _f32, _u64, _sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_uint64, C.c_size_t

# name -> (restype, argtypes); mirrors include/b200p.h one to one
SIGNATURES = {
    "b200p_version": (C.c_char_p, []),
    "b200p_last_error": (C.c_char_p, []),
    "b200p_arch_check": (_i32, []),
    "b200p_sm_count": (_i32, []),
    "b200p_set_pdl": (_i32, [_i32]),
    "b200p_dequantize": (_i32, [_vp, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200p_quantize": (_i32, [_vp, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _vp]),
    "b200p_embed": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200p_qmm": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _vp, _sz, _vp]),
    "b200p_qmm_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "b200p_qmm_tensor_supported": (_i32, [_i64, _i64, _i64, _i32, _i32, _i32, _i32]),
    "b200p_qmm_set_path": (_i32, [_i32]),
    "b200p_qmm_set_min_tensor_m": (_i32, [_i32]),
    "b200p_rmsnorm": (_i32, [_vp, _vp, _vp, _i64, _i64, _f32, _i32, _i32, _vp]),
    "b200p_rope_kv_append": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _i64,
                                    _f32, _f32, _f32, _i32, _i32, _vp]),
    "b200p_silu_mul": (_i32, [_vp, _vp, _i64, _i64, _i64, _i64, _i32, _i32, _vp]),
    "b200p_lora_add": (_i32, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64, _f32, _i32, _i32, _vp]),
    "b200p_advance_lengths": (_i32, [_vp, _i64, _i32, _vp]),
    "b200p_kv_overflow_count": (_i32, [C.POINTER(C.c_uint), _i32, _vp]),
    "b200p_attn": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _i64, _f32, _i32, _i32, _vp, _sz, _vp]),
    "b200p_attn_workspace_bytes": (_sz, [_i64, _i64, _i32, _i32]),
    "b200p_attn_decode_static": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i64, _f32, _i32, _i32, _vp]),
    "b200p_attn_set_path": (_i32, [_i32]),
    "b200p_argmax": (_i32, [_vp, _vp, _vp, _i64, _i64, _i64, _vp]),
    "b200p_argmax_partial": (_i32, [_vp, _vp, _i64, _i64, _i64, _i32, _vp]),
    "b200p_argmax_merge": (_i32, [_vp, _vp, _vp, _i32, _i64, _vp]),
    "b200p_sample_top_p": (_i32, [_vp, _vp, _vp, _i64, _i64, _i64, _f32, _f32, _u64, _u64, _vp]),
    "b200p_symm_alloc": (_i32, [C.POINTER(_vp), _sz]),
    "b200p_symm_free": (_i32, [_vp]),
    "b200p_ipc_export": (_i32, [_vp, _vp]),
    "b200p_ipc_import": (_i32, [_vp, C.POINTER(_vp)]),
    "b200p_ipc_close": (_i32, [_vp]),
    "b200p_allreduce_flag_bytes": (_sz, []),
    "b200p_allreduce_set_max_ctas": (_i32, [_i32]),
    "b200p_allreduce_residual_rmsnorm": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _f32, _i32, _i32, _vp]),
    "b200p_allreduce_rows_rmsnorm": (_i32, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _i64, _i64, _i64, _f32, _i32, _i32, _vp]),
    "b200p_peer_allgather": (_i32, [_vp, _vp, _i32, _i32, _vp, _i64, _vp]),
    "b200p_peer_status": (_i32, [_vp, _vp, _vp]),
}

_lib = None


class B200PError(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200PError(
                f"{LIB_PATH} not found: the CUDA extension is not built and there is no CPU fallback "
                "(run `make -C mlx_parallm_b200/csrc`)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError here = header/library drift
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def last_error() -> str:
    return lib().b200p_last_error().decode()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise B200PError(f"{what or 'b200p call'} failed ({rc}): {last_error()}")


def require_device() -> None:
    """Fail loudly unless an sm_100 device is current."""
    check(lib().b200p_arch_check(), "b200p_arch_check")
