"""ctypes binding of ``libobjcavit_b200.so`` (the C ABI declared in ``include/objcavit_b200.h``).

There is no fallback: if the library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libobjcavit_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "objcavit_b200.h")

_lib = None

p = C.c_void_p
i32 = C.c_int
i64 = C.c_longlong
f32 = C.c_float

# name -> (restype, argtypes): must mirror include/objcavit_b200.h
PROTOTYPES = {
    "ocv_abi_version": (i32, []),
    "ocv_device_check": (i32, [i32]),
    "ocv_set_dropout_salt": (i32, [p]),
    "ocv_release_dropout_salt": (i32, [p]),
    "ocv_status_string": (C.c_char_p, [i32]),
    "ocv_linear_f32": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, i32, i32, i32, i32, i64, i64, i64, p]),
    "ocv_patch_embed_f32_workspace": (i64, [i32, i32, i32, i32, i32, i32]),
    "ocv_patch_embed_f32": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, i32, p, p]),
    "ocv_add_layernorm_f32": (i32, [p, p, p, p, p, i32, i32, p, p]),
    "ocv_mha_core_f32": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32, i64, i32, p]),
    "ocv_pad_objects_f32": (i32, [p, p, p, p, p, i32, i32, i32, i32, f32, i32, p]),
    "ocv_patch_coords_f32": (i32, [p, i32, i32, i32, p]),
    "ocv_grid_sample_pos_f32": (i32, [p, p, p, p, i32, i32, i32, i32, i32, f32, f32, p]),
    "ocv_ps_roi_align_pos_f32": (i32, [p, p, p, p, i32, i32, i32, i32, f32, i32, p]),
    "ocv_bins_f32": (i32, [p, p, p, p, i32, i32, i32, f32, f32, p]),
    "ocv_conv3x3_f32": (i32, [p, p, p, p, i32, i32, i32, i32, p]),
    "ocv_tail_f32": (i32, [p, p, p, p, p, i32, i32, i32, i64, p]),
    "ocv_nchw_to_nhwc_bf16": (i32, [p, p, i32, i32, i32, i32, i32, p]),
    "ocv_gemm_bf16_workspace": (i64, [i32, i32, i32, i32]),
    "ocv_gemm_bf16": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, i32, i32, i32, i32, p, p]),
    "ocv_mha_core_tc_bf16": (i32, [p, p, p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32, i64, i32, f32,
                                    C.c_ulonglong, p]),
    "ocv_mha_core_bf16": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32, i64, i32, p]),
    "ocv_pack_patch_weights_bf16": (i32, [p, p, i32, i32, i32, p]),
    "ocv_patch_embed_bf16_workspace": (i64, [i32, i32, i32, i32]),
    "ocv_patch_embed_bf16": (i32, [p, p, p, p, p, p, i32, i32, i32, i32, i32, i32, p, p]),
    "ocv_mlp_f32": (i32, [p, i64, i32, i32, p, p, p, p, p, p, p]),
    "ocv_cast_f32_to_bf16": (i32, [p, p, i64, p]),
    "ocv_linear_rows_f32": (i32, [p, i64, i32, p, p, i32, i32, i32, p, i32, p]),
    "ocv_mlp_rows_f32_workspace": (i64, [i32, i32, p]),
    "ocv_mlp_rows_f32": (i32, [p, i64, i32, i32, p, p, p, p, p, p]),
    "ocv_gemm_bf16_v2": (i32, [p, p, i32, i32, p, p, i32, p, p, p, i32, p, i32, p, i32, i32, i32, i32, i32, i32, i32, f32, C.c_ulonglong,
                                i32, p, p]),
    "ocv_gemm_bf16_ex": (i32, [p, p, p, p, i32, p, p, p, p, i32, p, i32, i32, i32, i32, i32, i32, i32, p]),
    "ocv_pack_conv3x3_bf16": (i32, [p, p, p]),
    "ocv_fold_queries_bf16": (i32, [p, p, p, i64, i32, p, f32, p, p, i32, i32, i32, i32, p]),
    "ocv_pixel_fused_bf16": (i32, [p, p, p, p, p, p, i32, i32, i32, i32, p]),
    "ocv_conv3x3_nhwc_bf16": (i32, [p, p, p, p, i32, i32, i32, i32, p]),
    "ocv_pixel_fused_bf16_argmax": (i32, [p, p, p, p, p, p, p, i32, i32, i32, p]),
    "ocv_linear_splitk_f32": (i32, [p, p, p, p, i32, i32, i32, i32, i32, i32, i32, p, p]),
    "ocv_gemm_f32_ex_workspace": (i64, [i32, i32, i32, i32]),
    "ocv_gemm_f32_ex": (i32, [i32, i32, p, p, p, i32, i32, i32, i32, i32, i32, f32, i32, i64, i64, i64, i32, i32, p, p, p]),
    "ocv_reduce_channels_f32": (i32, [p, p, i32, i32, i32, i64, f32, p]),
    "ocv_act_bwd_f32": (i32, [p, p, p, i64, i32, p]),
    "ocv_layernorm_bwd_f32": (i32, [p, p, p, p, p, p, i32, i32, f32, p, p]),
    "ocv_layernorm_bwd_noalloc_f32": (i32, [p, p, p, p, p, p, i32, i32, f32, p]),
    "ocv_layernorm_bwd_params_f32": (i32, [p, p, p, p, i32, i32, f32, p, p]),
    "ocv_layernorm_bwd_dropout_f32": (i32, [p, p, p, p, p, i32, i32, f32, C.c_ulonglong, p, p]),
    "ocv_mha_core_bwd_f32": (i32, [p, p, p, p, p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32, i64, i32,
                                   i64, i32, i64, i32, i64, i32, p, p]),
    "ocv_tail_bwd_f32": (i32, [p, p, p, p, p, i32, i32, i64, p]),
    "ocv_dropout_f32": (i32, [p, p, p, i64, f32, C.c_ulonglong, p]),
    "ocv_dropout_add_layernorm_f32": (i32, [p, p, p, p, p, p, i32, i32, f32, C.c_ulonglong, p]),
    "ocv_dropout_act_bwd_f32": (i32, [p, p, p, i64, i32, f32, C.c_ulonglong, p]),
    "ocv_encoder_layer_fwd_train_workspace": (i64, [i32, i32, i32]),
    "ocv_encoder_layer_fwd_train_f32": (i32, [p, p, p, i32, i32, i32, f32, p, p, p, p, p, p, p, p, p, p]),
    "ocv_encoder_layer_bwd_train_workspace": (i64, [i32, i32, i32]),
    "ocv_encoder_layer_fwd_train_bf16": (i32, [p, p, p, p, p, i32, i32, i32, f32, p, p, p, p, p, p, p, p, p, p, p, p, p, p, p]),
    "ocv_mha_core_bwd_tc_bf16": (i32, [p, p, p, p, p, p, p, p, p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32,
                                        i64, i32, f32, C.c_ulonglong, p]),
    "ocv_encoder_layer_bwd_train_bf16_workspace": (i64, [i32, i32, i32]),
    "ocv_encoder_layer_bwd_train_bf16": (i32, [p, p, p, p, p, p, p, p, p, p, p, p, p, p, p, p, p, i32, i32, i32, f32, p, p, p, p]),
    "ocv_encoder_layer_bwd_train_f32": (i32, [p, p, p, p, p, p, p, p, p, p, p, p, i32, i32, i32, f32, p, p, p, p]),
    "ocv_grid_sample_pos_bwd_f32": (i32, [p, p, p, i32, i32, i32, i32, i32, f32, f32, p]),
    "ocv_ps_roi_align_pos_bwd_f32": (i32, [p, p, p, i32, i32, i32, i32, f32, i32, p]),
    "ocv_table_grad_commit_f32": (i32, [p, p, i64, f32, p]),
    "ocv_tail_fused_fwd_bf16": (i32, [p, p, p, p, p, p, p, i32, i64, p]),
    "ocv_tail_fused_bwd_workspace": (i64, [i32, i64]),
    "ocv_tail_fused_bwd_bf16": (i32, [p, p, p, p, p, p, p, p, p, i64, p, p, p, f32, i32, i64, p, p]),
    "ocv_tail_logits_f32": (i32, [p, p, p, p, i32, i32, i64, p]),
    "ocv_tail_logits_bwd_workspace": (i64, [i32, i32, i64]),
    "ocv_tail_logits_stats_f32": (i32, [p, p, p, p, p, i32, i32, i64, p]),
    "ocv_tail_logits_bwd_stats_bf16": (i32, [p, p, p, p, p, p, p, p, f32, i32, i32, i64, p, p]),
    "ocv_tail_logits_bwd_bf16": (i32, [p, p, p, p, p, p, p, f32, i32, i32, i64, p, p]),
    "ocv_transpose_bf16": (i32, [p, p, i32, i32, i32, i64, i64, i64, i64, p]),
    "ocv_im2col3x3_bf16": (i32, [p, p, i32, i32, i32, i32, p]),
    "ocv_patch_unfold_bf16": (i32, [p, p, i32, i32, i32, i32, i32, i64, p]),
    "ocv_patch_unfold_nhwc_bf16": (i32, [p, p, i32, i32, i32, i32, i32, i64, p]),
    "ocv_conv3x3_wgrad_bf16_workspace": (i64, [i32, i32, i32]),
    "ocv_conv3x3_wgrad_bf16": (i32, [p, i64, p, p, p, i32, i32, i32, f32, p, p]),
    "ocv_transpose_cast_f32_bf16": (i32, [p, p, i32, i32, i64, i64, p]),
    "ocv_mha_core_dropout_f32": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32, i64, i32, f32,
                                       C.c_ulonglong, p]),
    "ocv_mha_core_bwd_dropout_f32": (i32, [p, p, p, p, p, p, p, p, p, i32, i32, i32, i32, i32, i64, i32, i64, i32, i64, i32, i64, i32,
                                           i64, i32, i64, i32, i64, i32, f32, C.c_ulonglong, p, p]),
    "ocv_unpad_objects_f32": (i32, [p, p, p, i32, i32, i32, i32, i32, f32, p]),
    "ocv_tta_merge_f32": (i32, [p, p, p, i32, i32, i32, f32, f32, p]),
    "ocv_depth_metrics_workspace": (i64, []),
    "ocv_depth_metrics_f32": (i32, [p, p, i32, i32, i32, i32, i32, f32, f32, p, p, p, p, p, p]),
    "ocv_object_glue_f32": (i32, [p, p, p, p, i32, i32, i32, i32, p]),
    "ocv_sumsq_workspace": (i64, [i64]),
    "ocv_sumsq_f32": (i32, [p, i64, f32, p, p, p]),
    "ocv_adamw_step_f32": (i32, [p, p, p, p, i64, f32, f32, f32, f32, f32, i32, f32, p, f32, p]),
    "ocv_bins_bwd_f32": (i32, [p, p, p, p, i32, i32, f32, f32, p]),
    "ocv_bins_bwd_ex_f32": (i32, [p, p, p, p, p, i32, i32, f32, f32, p]),
    "ocv_silog_workspace": (i64, [i32, i32, i32]),
    "ocv_silog_forward": (i32, [p, p, p, i32, i32, i32, i32, i32, f32, f32, p, p, p]),
    "ocv_silog_backward": (i32, [p, p, p, i32, i32, i32, i32, i32, f32, f32, f32, p, p, p]),
    "ocv_bins_chamfer_workspace": (i64, [i32, i32]),
    "ocv_bins_chamfer": (i32, [p, p, p, i32, i32, i64, f32, p, p, p, p]),
    "ocv_head_create": (i32, [p, p]),
    "ocv_head_destroy": (None, [p]),
    "ocv_head_set_param": (i32, [p, C.c_char_p, p, i64]),
    "ocv_head_packed_bytes": (i64, [p]),
    "ocv_head_pack": (i32, [p, p, p]),
    "ocv_head_missing_key": (C.c_char_p, [p]),
    "ocv_head_workspace_bytes": (i64, [p, i32, i32, i32, i32, i32]),
    "ocv_head_forward": (i32, [p, p, p, p, p, i32, i32, i32, i32, i32, p, p, p, p, p]),
    "ocv_head_last_launches": (i32, [p]),
    "ocv_head_set_pipeline": (i32, [p, i32, i32]),
    "ocv_head_join": (i32, [p, p]),
    "ocv_head_stream": (p, [p, i32, i32]),
    "ocv_head_enable_timing": (i32, [p, i32]),
    "ocv_head_pixel_ms": (f32, [p]),
}


class HeadConfig(C.Structure):
    """``ocv_head_config`` of include/objcavit_b200.h."""
    _fields_ = [("kind", i32), ("strategy", i32), ("no_obj_sa", i32), ("use_2_saca", i32), ("n_bins", i32),
                ("obj_dim", i32), ("min_depth", f32), ("max_depth", f32), ("reserved", i32 * 8)]


def declared_symbols():
    """Every ``ocv_*`` function the public header declares (used by the symbol-export test)."""
    with open(HEADER_PATH) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ocv_[a-z0-9_]+)\s*\(", src)))


class OcvError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OcvError(
                f"{LIB_PATH} is missing: build it with `python -m objcavit_b200.build` "
                "(or __graft_entry__.build()). There is no CPU / PyTorch fallback for the kernels.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the .so is stale -> loud
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status: int, what: str = ""):
    if status != 0:
        msg = lib().ocv_status_string(status).decode()
        raise OcvError(f"objcavit_b200 {what} failed: status {status} ({msg})")
