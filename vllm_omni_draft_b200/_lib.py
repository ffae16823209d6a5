"""ctypes binding of libqi_b200.so (the C ABI declared in include/qi_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``csrc/Makefile``.  There is no
fallback: if the shared object is missing, or a compute entry point reports an error, a
``RuntimeError`` is raised (reference error convention: RuntimeError for load failures,
src/qwen_image_gen/model.py:72-75)."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
# QI_LIB_PATH: A/B aid (a differently built copy of the same library); the default is the in-tree build
LIB_PATH = os.environ.get("QI_LIB_PATH") or os.path.join(_HERE, "libqi_b200.so")

_lib: Optional[C.CDLL] = None

_vp, _i, _ll, _f = C.c_void_p, C.c_int, C.c_longlong, C.c_float


class GemmProblem(C.Structure):
    """qi_gemm_problem_t"""
    _fields_ = [
        ("a", _vp), ("lda", _ll), ("a_batch_stride", _ll),
        ("w", _vp), ("bias", _vp),
        ("c", _vp), ("ldc", _ll), ("c_batch_stride", _ll),
        ("residual", _vp), ("ld_res", _ll), ("res_batch_stride", _ll),
        ("gate", _vp), ("gate_batch_stride", _ll),
        ("m", _i), ("batch", _i), ("n", _i), ("k", _i), ("epilogue", _i),
        ("a_k_chunk", _i), ("a_k_chunk_stride", _ll),
        ("qk_norm_wq", _vp), ("qk_norm_wk", _vp), ("rope_cos_sin", _vp),
        ("rope_pos0", _i), ("heads", _i), ("head_dim", _i), ("qk_norm_eps", C.c_float),
        ("scatter", _vp),
        ("c_fp32", _i), ("gate_fp32", _i),
        ("wait", _vp),
    ]


class NormRows(C.Structure):
    """qi_norm_rows_t"""
    _fields_ = [("x", _vp), ("y", _vp), ("rows", _i), ("ld_x", _i), ("ld_y", _i), ("shift", _vp), ("scale", _vp),
                ("mod_batch_stride", _ll), ("rows_per_batch", _i), ("x_fp32", _i), ("mod_fp32", _i)]


class PeerSignal(C.Structure):
    """qi_peer_signal_t"""
    _fields_ = [("slots", _vp), ("state", _vp), ("n_peers", _i), ("my_index", _i), ("timeout_ms", C.c_uint)]


class PeerScatter(C.Structure):
    """qi_peer_scatter_t"""
    _fields_ = [("n_peers", _i), ("my_index", _i), ("base", _vp * 8), ("row_prefix", _i * 9),
                ("chunk_stride", _ll), ("notify", _vp), ("wait", _vp)]


_SIGNATURES = {
    "qi_version": (_i, []),
    "qi_last_cuda_error": (C.c_char_p, []),
    "qi_device_sm_count": (_i, []),
    "qi_set_pdl": (_i, [_i]),
    "qi_gemm_bf16": (_i, [C.POINTER(GemmProblem), _i, _vp]),
    "qi_gemm_set_tile_n": (_i, [_i]),
    "qi_gemm_set_cta_pair": (_i, [_i]),
    "qi_attention_set_variant": (_i, [_i]),
    "qi_gemm_set_debug": (_i, [_i]),
    "qi_gemm_set_schedule": (_i, [_vp, _i, _i, _i]),
    "qi_gemm_clear_schedules": (_i, []),
    "qi_gemm_set_group_m": (_i, [_i]),
    "qi_gemm_set_schedule_ex": (_i, [_vp, _i, _i, _i, _i]),
    "qi_attention_fwd_scatter": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _ll, _ll, _ll, _ll, _ll, _ll, _ll, _ll, _ll,
                                      _f, C.POINTER(PeerScatter), _vp]),
    "qi_qknorm_rope_scatter": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f,
                                    C.POINTER(PeerScatter), _vp]),
    "qi_peer_alloc": (_i, [C.POINTER(_vp), C.c_ulonglong]),
    "qi_peer_free": (_i, [_vp]),
    "qi_peer_export": (_i, [_vp, _vp]),
    "qi_peer_import": (_i, [_vp, C.POINTER(_vp)]),
    "qi_peer_unmap": (_i, [_vp]),
    "qi_peer_barrier": (_i, [_vp, _i, _i, _vp, C.c_uint, _vp]),
    "qi_peer_barrier_static": (_i, [C.POINTER(_vp), _i, _i, _i, _vp, C.c_uint, _vp]),
    "qi_attention_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _ll, _ll, _ll, _ll, _ll, _ll, _f, _vp]),
    "qi_adaln_modulate": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _ll, _i, _f, _vp]),
    "qi_adaln_modulate_multi": (_i, [C.POINTER(NormRows), _i, _i, _f, _vp]),
    "qi_rmsnorm_affine": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _f, _vp]),
    "qi_qknorm_rope": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp]),
    "qi_cfg_euler_step": (_i, [_vp, _vp, _vp, _vp, _ll, _i, _f, _vp, _f, _vp]),
    "qi_cfg_euler_step_f32": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _i, _f, _vp, _f, _vp]),
    "qi_small_linear": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "qi_timestep_sinusoid": (_i, [_vp, _vp, _i, _vp]),
    "qi_small_linear_f32": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "qi_timestep_sinusoid_f32": (_i, [_vp, _vp, _i, _vp]),
    "qi_rope_position_ids": (_i, [C.POINTER(C.c_int), _i, _i, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "qi_pack_latents_index": (_i, [_i, _i, _i, C.POINTER(C.c_int64)]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_STATUS = {-1: "invalid argument", -2: "CUDA error / no sm_100 device", -3: "tensor-map encode failed",
           -4: "unsupported shape"}


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU or PyTorch fallback for the hot path.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        detail = lib().qi_last_cuda_error()
        raise RuntimeError(f"{what} failed: {_STATUS.get(rc, rc)}"
                           + (f" ({detail.decode()})" if detail else ""))
