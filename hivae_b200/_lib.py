"""ctypes binding of libhivae_b200.so (the C ABI declared in include/hivae_b200.h).

There is no CPU fallback: every op raises if the shared library is missing or the
device is not an sm_100 GPU.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhivae_b200.so")

HV_F32, HV_BF16 = 0, 1
EPI_STORE, EPI_GELU, EPI_RESIDUAL, EPI_QKNORM = 0, 1, 2, 3

_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float


class HvEpilogue(C.Structure):
    _fields_ = [
        ("mode", _i), ("out_dtype", _i), ("out", _vp), ("ldo", _i64), ("bias", _vp),
        ("res", _vp), ("L", _i), ("frame_rows", _i64), ("row_off", _i),
        ("gate", _vp), ("gate_ld", _i64), ("gate_col_mot", _i), ("gate_col_img", _i),
        ("img_start", _i), ("img_len", _i), ("cond_div", _i), ("cond_mul", _i),
        ("qn_w", _vp), ("qn_b", _vp), ("kn_w", _vp), ("kn_b", _vp),
        ("q_cols", _i), ("k_cols", _i), ("qk_eps", _f), ("k_split", _i), ("out2", _vp), ("ldo2", _i64),
    ]


class HvSeqMap(C.Structure):
    _fields_ = [("sdiv", _i), ("sa", _i64), ("sb", _i64), ("sc", _i64)]


# name -> (restype, argtypes); mirrors include/hivae_b200.h one to one
SIGNATURES = {
    "hv_last_error": (C.c_char_p, []),
    "hv_version": (_i, []),
    "hv_check_device": (_i, []),
    "hv_launch_count": (_i64, []),
    "hv_dwt2_haar": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "hv_dwt2_haar_pyramid": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "hv_idwt2_haar_pyramid": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "hv_idwt2_haar": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "hv_freq_split3d": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "hv_ln_modulate": (_i, [_vp, _i64, _i64, _i, _i, _i, _vp, _vp, _f, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "hv_ln_modulate_save": (_i, [_vp, _i64, _i64, _i, _i, _i, _vp, _vp, _f, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp]),
    "hv_patch_embed": (_i, [_vp, _i, _vp, _i, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _i, _i64, _i, _i, _vp]),
    "hv_final_head": (_i, [_vp, _i64, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _f, _vp, _i64, _i, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _f, _vp]),
    "hv_timestep_embed": (_i, [_vp, _i64, _i, _vp, _i, _vp]),
    "hv_silu": (_i, [_vp, _i, _vp, _i, _i64, _vp]),
    "hv_convert": (_i, [_vp, _i, _vp, _i, _i64, _vp]),
    "hv_axpy": (_i, [_f, _vp, _vp, _i64, _vp]),
    "hv_rows_scatter": (_i, [_vp, _i64, _i64, _vp, _vp, _i64, _i64, _i, _i, _i, _vp]),
    "hv_rows_gather": (_i, [_vp, _i64, _i64, _i, _i, _i, _vp, _vp]),
    "hv_attn_prof_read": (_i, [_vp]),
    "hv_attn_timeline_read": (_i, [_vp]),
    "hv_attn_bwd_prof_read": (_i, [_vp]),
    "hv_headnorm64_bwd": (_i, [_vp, _i64, _vp, _i64, _vp, _f, _i64, _i, _vp, _i64, _i, _vp, _vp, _vp, _vp]),
    "hv_headnorm64_bwd_acc": (_i, [_vp, _i64, _vp, _i64, _vp, _f, _i64, _i, _vp, _i64, _i, _vp, _vp, _vp, _i, _vp]),
    "hv_attention_bwd": (_i, [_vp, _i64] * 8 + [_i, _i64, _i, _i, _i, _i, _vp, _vp, _f, _vp, _vp]),
    "hv_attention_bwd_lse": (_i, [_vp, _i64] * 8 + [_i, _i64, _i, _i, _i, _i, _vp, _vp, _f, _vp, _i, _vp]),
    "hv_transpose2d": (_i, [_vp, _vp, _i64, _i64, _i64, _i64, _i, _vp]),
    "hv_transpose2d_colsum": (_i, [_vp, _vp, _i64, _i64, _i64, _i64, _i, _vp, _vp, _vp]),
    "hv_colsum_slices": (_i, []),
    "hv_colsum": (_i, [_vp, _i64, _i, _i64, _i, _vp, _vp, _vp]),
    "hv_colsum_acc": (_i, [_vp, _i64, _i, _i64, _i, _vp, _vp, _i, _vp, _i, _vp, _i, _vp]),
    "hv_gelu_tanh_bwd": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "hv_layernorm_bwd_ctas": (_i, []),
    "hv_layernorm_bwd": (_i, [_vp, _vp, _i, _vp, _f, _i64, _i, _vp, _i, _vp, _vp, _vp, _vp]),
    "hv_unpatchify": (_i, [_vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp]),
    "hv_conv3x3": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp]),
    "hv_maxpool2": (_i, [_vp, _vp, _i64, _i, _i, _vp]),
    "hv_rope": (_i, [_vp, _vp, _i64, _i, _i, _i, _i64, _i, _f, _vp]),
    "hv_linear_f32": (_i, [_vp, _i, _i64, _vp, _i64, _i64, _i, _i, C.POINTER(HvEpilogue), _vp]),
    "hv_linear_bf16": (_i, [_vp, _i64, _vp, _i64, _i64, _i, _i, C.POINTER(HvEpilogue), _vp]),
    "hv_linear_bf16_tn": (_i, [_vp, _i64, _vp, _i64, _i64, _i, _i64, C.POINTER(HvEpilogue), _vp]),
    "hv_reduce_workspace_doubles": (_i, []),
    "hv_mse_loss": (_i, [_vp, _vp, _i64, _vp, _vp, _vp]),
    "hv_mse_loss_static": (_i, [_vp, _vp, _i64, _vp, _vp]),
    "hv_mse_loss_static_max_elements": (_i, []),
    "hv_grad_norm_clip_coef": (_i, [_vp, _i64, _f, _f, _vp, _vp, _vp]),
    "hv_adamw_clip": (_i, [_vp, _vp, _vp, _vp, _i64, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, _i64, _f, _vp, _vp]),
    "hv_adamw_clip_devstate": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, C.c_double, C.c_double, C.c_double, C.c_double, _f, _vp, _vp]),
    "hv_scale_by_device": (_i, [_vp, _i64, _vp, _f, _vp]),
    "hv_train_chunk_rows": (_i, []),
    "hv_gate_bwd": (_i, [_vp, _i64, _i64, _i, _i, _i, _vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _i64, _vp]),
    "hv_ln_modulate_bwd": (_i, [_vp, _vp, _i, _vp, _vp, _f, _vp, _i64, _i64, _i, _i, _i, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _i,
                                _vp, _vp, _vp, _vp, _i64, _vp]),
    "hv_gate_residual": (_i, [_vp, _i64, _i64, _i, _i, _i, _vp, _i, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp]),
    "hv_silu_bwd": (_i, [_vp, _vp, _i, _vp, _i, _i64, _vp]),
    "hv_mse_bwd": (_i, [_vp, _vp, _i64, _vp, _vp, _vp]),
    "hv_rows_reduce": (_i, [_vp, _i, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _vp, _i, _vp]),
    "hv_rows_index": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _i64, _i64, _i, _vp]),
    "hv_conv3x3_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp]),
    "hv_maxpool2_bwd": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "hv_attention": (_i, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _i, _i64, _i, _i, _i, _i,
                          C.POINTER(HvSeqMap), C.POINTER(HvSeqMap), _f, _vp, _vp]),
    "hv_attention_lse": (_i, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _i, _i64, _i, _i, _i, _i,
                              C.POINTER(HvSeqMap), C.POINTER(HvSeqMap), _f, _vp, _vp, _vp]),
}

_lib = None


def load():
    """dlopen the shared library and bind every declared symbol (raises if one is missing)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "hivae_b200: %s not found - build it with `python -m hivae_b200.build` "
                "(there is no CPU/eager fallback)" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise RuntimeError("hivae_b200 kernel call failed (%d): %s" % (rc, load().hv_last_error().decode()))


_device_ok = False


def require_device():
    """Fail loudly unless a B200-class (sm_100) device is current."""
    global _device_ok
    if not _device_ok:
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("hivae_b200 needs a CUDA device (sm_100a); no CPU fallback exists")
        check(load().hv_check_device())
        _device_ok = True
