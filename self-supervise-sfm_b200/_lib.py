"""ctypes binding of libsailrecon_b200.so — the C-ABI declared in include/sailrecon_b200.h.

There is deliberately no fallback: if the shared library is missing, or a compute entry point is called on a
machine without an sm_100 device, a RuntimeError is raised.  Signatures carry only plain C types (device
pointers as void*, sizes as int64, cudaStream_t as void*).
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

SR_OK, SR_ERR_SHAPE, SR_ERR_ARCH, SR_ERR_CUDA = 0, 1, 2, 3
SR_EPI_BIAS, SR_EPI_BIAS_GELU, SR_EPI_LS_RESIDUAL, SR_EPI_QKV, SR_EPI_BIAS_SILU = 0, 1, 2, 3, 4

_vp, _i64, _i32, _f32 = C.c_void_p, C.c_int64, C.c_int32, C.c_float


class SrEpilogue(C.Structure):
    """struct sr_epilogue (include/sailrecon_b200.h)."""

    _fields_ = [
        ("kind", _i32), ("out_bf16", _i32), ("out", _vp), ("ldo", _i64),
        ("bias", _vp), ("gamma", _vp), ("resid", _vp), ("ldr", _i64),
        ("qkv_dim", _i64), ("col_offset", _i64), ("head_dim", _i32), ("qk_norm", _i32),
        ("qn_w", _vp), ("qn_b", _vp), ("kn_w", _vp), ("kn_b", _vp), ("qk_eps", _f32),
        ("rope", _i32), ("rope_cos", _vp), ("rope_sin", _vp), ("rope_max_pos", _i32), ("pos", _vp),
        ("exact_residual", _i32),
    ]


class SrAttn(C.Structure):
    """struct sr_attn (include/sailrecon_b200.h)."""

    _fields_ = [
        ("q", _vp), ("ldq", _i64),
        ("k", _vp), ("v", _vp), ("ldkv", _i64),
        ("pk", _vp), ("pv", _vp), ("ldp", _i64), ("prefix_len", _i64),
        ("out", _vp), ("ldo", _i64),
        ("num_heads", _i32), ("head_dim", _i32), ("batch", _i32),
        ("q_len", _i64), ("q_batch_stride", _i64), ("kv_len", _i64), ("kv_batch_stride", _i64),
        ("scale", _f32),
        ("carry_m", _vp), ("carry_l", _vp), ("carry_acc", _vp), ("carry_in", _i32), ("carry_out", _i32),
        ("lse", _vp), ("lse_pad", _i64),
        ("variant", _i32),
        ("workspace", _vp), ("workspace_bytes", _i64),
    ]


class SrAttnBwd(C.Structure):
    """struct sr_attn_bwd (include/sailrecon_b200.h)."""

    _fields_ = [
        ("q", _vp), ("ldq", _i64),
        ("k", _vp), ("v", _vp), ("ldkv", _i64),
        ("pk", _vp), ("pv", _vp), ("ldp", _i64), ("prefix_len", _i64),
        ("out", _vp), ("ldo", _i64),
        ("dout", _vp), ("lddo", _i64),
        ("lse", _vp), ("delta", _vp), ("lse_pad", _i64),
        ("dq", _vp), ("lddq", _i64),
        ("dk", _vp), ("dv", _vp), ("lddkv", _i64),
        ("dpk", _vp), ("dpv", _vp), ("lddp", _i64),
        ("num_heads", _i32), ("head_dim", _i32), ("batch", _i32),
        ("q_len", _i64), ("q_batch_stride", _i64), ("kv_len", _i64), ("kv_batch_stride", _i64),
        ("scale", _f32),
    ]


class SrReprojLoss(C.Structure):
    """struct sr_reproj_loss_t (include/sailrecon_b200.h)."""

    _fields_ = [
        ("pred_intrinsics", _vp), ("pred_extrinsics", _vp), ("kprime_to_k", _vp),
        ("src_idx", _vp), ("dst_idx", _vp),
        ("src_coords", _vp), ("dst_coords", _vp), ("src_depth", _vp), ("dst_depth", _vp),
        ("num_frames", _i64), ("num_pairs", _i64), ("num_points", _i64),
        ("pose_rows", _i32), ("shared_focal", _i32),
        ("min_val", _f32), ("inv_bin_width", _f32), ("sobel_weight", _f32),
        ("num_bins", _i32), ("smooth_radius", _i32),
        ("smooth_kernel", _vp),
        ("workspace", _vp), ("workspace_floats", _i64),
        ("loss", _vp), ("d_intrinsics", _vp), ("d_extrinsics", _vp),
        ("tables", _vp),
    ]


# name -> (restype, argtypes); every symbol include/sailrecon_b200.h declares
SIGNATURES = {
    "sr_version": (C.c_int, []),
    "sr_last_error": (C.c_char_p, []),
    "sr_arch_check": (C.c_int, []),
    "sr_num_kernel_launches": (C.c_int, []),
    "sr_layernorm_fwd": (C.c_int, [_vp, _i64, _vp, _vp, _vp, C.c_int, _i64, _i64, _i64, _f32, _vp]),
    "sr_layernorm_gather_fwd": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, C.c_int, _i64, _i64, _i64, _f32, _vp]),
    "sr_gemm_bf16": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _i64, C.POINTER(SrEpilogue), _vp]),
    "sr_gemm_bf16_t": (C.c_int, [_vp, _i64, _i32, _vp, _i64, _i32, _i64, _i64, _i64, C.POINTER(SrEpilogue), _vp]),
    "sr_gemm_f32": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _i64, C.POINTER(SrEpilogue), _vp]),
    "sr_debug_trace_read": (C.c_int, [C.POINTER(C.c_longlong), _i64]),
    "sr_attn_fwd_bf16": (C.c_int, [C.POINTER(SrAttn), _vp]),
    "sr_attn_fwd_f32": (C.c_int, [C.POINTER(SrAttn), _vp]),
    "sr_attn_workspace_bytes": (_i64, [C.POINTER(SrAttn), _i32]),
    "sr_attn_plan": (C.c_int, [_i64, _i32, _i32, _i32, C.POINTER(_i32), C.POINTER(_i32)]),
    "sr_attn_bwd_bf16": (C.c_int, [C.POINTER(SrAttnBwd), _vp]),
    "sr_transpose_bf16": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _i64, _vp, _vp]),
    "sr_colsum": (C.c_int, [_vp, C.c_int, _i64, _vp, _i64, _vp, _i64, _i64, C.c_int, _vp]),
    "sr_scale_cast_bf16": (C.c_int, [_vp, _i64, _vp, _vp, _i64, _i64, _i64, _vp]),
    "sr_ls_residual_f32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _vp]),
    "sr_gelu_fwd_bf16": (C.c_int, [_vp, _vp, _i64, _vp]),
    "sr_gelu_bwd_bf16": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
    "sr_layernorm_bwd": (C.c_int, [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _i64, _i64, _f32, _vp]),
    "sr_qknorm_rope_fwd": (C.c_int, [_vp, _i64, _vp, _i64, C.POINTER(SrEpilogue), _i64, _vp]),
    "sr_qknorm_rope_bwd": (C.c_int, [_vp, _i64, _vp, _i64, C.POINTER(SrEpilogue), _vp, _vp, _vp, _vp, _i64, _vp]),
    "sr_scatter_add_rows_f32": (C.c_int, [_vp, _i64, _vp, _vp, _i64, _i64, _i64, _vp]),
    "sr_adam_step_f32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, C.c_double, C.c_double, C.c_double, C.c_double, _i64, C.c_double, _vp]),
    "sr_gather_rows_f32": (C.c_int, [_vp, _i64, _vp, _vp, _i64, _i64, _i64, _vp]),
    "sr_copy_cols_f32": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _i64, _vp]),
    "sr_cast_f32_to_bf16": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _vp]),
    "sr_adaln_modulate_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _f32, _vp]),
    "sr_scale_cols_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _vp]),
    "sr_rope_table_build": (C.c_int, [_vp, _vp, _i32, _i32, _f32, _vp]),
    "sr_rope2d_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i64, _i64, _i64, _i64, _vp]),
    "sr_patchify": (C.c_int, [_vp, _vp, C.c_int, _i64, _i64, _i64, _i64, _i32, C.POINTER(_f32), C.POINTER(_f32), _vp]),
    "sr_pose_update": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp]),
    "sr_pose_to_extri_intri": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _i32, _vp]),
    "sr_pose_to_extri_intri_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _vp]),
    "sr_unproject_depth": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp]),
    "sr_im2col_split_bf16": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _vp]),
    "sr_bilinear_nhwc_f32": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64, _i32, _vp]),
    "sr_conv3x3_bf16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _i64, C.POINTER(SrEpilogue), _vp]),
    "sr_split_pad_bf16": (C.c_int, [_vp, _i64, _i32, _vp, _i64, _i64, _i64, _i64, _i32, _vp]),
    "sr_map_add_f32": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _i32, _i64, _i64, _i64, _i64, _i32, _vp]),
    "sr_pixel_shuffle_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _i32, _vp]),
    "sr_add_bcast_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _f32, _i32, _vp]),
    "sr_uv_posembed_table": (C.c_int, [_vp, _i64, _i64, _i64, C.c_double, C.c_double, _vp]),
    "sr_dpt_out_head_f32": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "sr_modulate_f32": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _i64, _i64, _vp]),
    "sr_relative_pose_f32": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _vp]),
    "sr_relative_pose_bwd_f32": (C.c_int, [_vp, _vp, _i32, _vp, _vp, _vp, _i64, _vp]),
    "sr_reproject_workspace_floats": (_i64, [_i64, _i64]),
    "sr_reproject_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "sr_reproject_static_max_pairs": (_i64, []),
    "sr_reproject_static_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i64, _i64, _vp]),
    "sr_reproject_bwd_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "sr_projective_residual_f32": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
    "sr_projective_residual_bwd_f32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp]),
    "sr_cdf_tables_f32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _f32, _f32, _i32, _f32, _vp, _i32, _vp, _vp, _vp, _vp, _vp]),
    "sr_cdf_lookup_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _f32, _f32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "sr_reproj_loss_workspace_floats": (_i64, [_i64, _i64, _i64, _i32]),
    "sr_reproj_cdf_loss_f32": (C.c_int, [C.POINTER(SrReprojLoss), _vp]),
    "sr_ls_bwd_bf16": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _vp, _i64, _i64, _vp]),
    "sr_gelu_bwd_colsum_bf16": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _i64, _vp]),
    "sr_transpose_f32": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i64, _vp]),
    "sr_colsum_f32": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _i64, _i64, _i32, _vp]),
    "sr_act_f32": (C.c_int, [_vp, _vp, _i64, _i32, _vp]),
    "sr_act_bwd_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp]),
    "sr_ls_residual_f32f32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _vp]),
    "sr_layernorm_bwd_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _f32, _vp]),
    "sr_adaln_modulate_bwd_f32": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _f32, _i32, _vp]),
    "sr_attn_small_bwd_f32": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i64, _f32, _vp]),
    "sr_pose_act_bwd_f32": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
}

_lib = None


def lib_path() -> str:
    return os.environ.get("SAILRECON_B200_LIB", _build.LIB_PATH)


def load():
    """dlopen the library (building it first if nvcc is available and it is stale) and bind every symbol."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if path == _build.LIB_PATH and not _build.is_current():
        try:
            _build.build_library()                 # compile / link errors propagate: never run stale kernels silently
        except _build.NvccNotFound as e:           # no compiler here: a prebuilt library may be used, loudly
            if not os.path.exists(path):
                raise RuntimeError(
                    f"libsailrecon_b200.so is not built and could not be built ({e}); "
                    "sailrecon_b200 has no CPU or PyTorch fallback") from e
            import warnings
            warnings.warn(f"sailrecon_b200: {path} does not match the digest of csrc/ + the header and nvcc is not available "
                          "to rebuild it; loading the STALE library", RuntimeWarning)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} not found; sailrecon_b200 has no CPU or PyTorch fallback")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc == SR_OK:
        return
    msg = load().sr_last_error().decode(errors="replace")
    kind = {SR_ERR_SHAPE: "bad argument", SR_ERR_ARCH: "unsupported device", SR_ERR_CUDA: "CUDA error"}.get(rc, "error")
    if rc == SR_ERR_SHAPE:
        raise ValueError(f"sailrecon_b200 {what}: {kind}: {msg}")
    raise RuntimeError(f"sailrecon_b200 {what}: {kind}: {msg}")
