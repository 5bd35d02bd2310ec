"""ctypes binding of the C-ABI library (include/avion_b200.h -> avion_b200/csrc/libavion_b200.so).

There is no fallback of any kind: if the library is missing or a call fails, this raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# AVN_LIB_PATH: load another build of the same C-ABI (kernel A/B comparisons inside one GPU call; tools/ only)
LIB_PATH = os.environ.get("AVN_LIB_PATH") or os.path.join(_HERE, "csrc", "libavion_b200.so")

GEMM_NT, GEMM_NN, GEMM_TN = 0, 1, 2
EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_GELU_DUAL, EPI_BIAS_RESIDUAL, EPI_DGELU, EPI_F32_ADD = range(7)

_vp, _i, _i64, _f, _d, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_uint64


class GemmDesc(C.Structure):
    _fields_ = [("mode", _i), ("epi", _i), ("M", _i), ("N", _i), ("K", _i), ("batch", _i),
                ("A", _vp), ("lda", _i64), ("a_batch_stride", _i64),
                ("B", _vp), ("ldb", _i64), ("b_batch_stride", _i64),
                ("C", _vp), ("ldc", _i64), ("c_batch_stride", _i64),
                ("C2", _vp), ("bias", _vp),
                ("R", _vp), ("ldr", _i64), ("r_batch_stride", _i64),
                ("splits", _i), ("colsum", _vp)]


# name -> argtypes; every function returns int (0 = ok) unless noted
SIGNATURES = {
    "avn_version": [],
    "avn_device_check": [],
    "avn_sample_rrc_boxes": [_u64, _i, _i, _i, _d, _d, _d, _d, _d, _vp],
    "avn_clip_transform_u8_bf16": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, C.POINTER(_f), C.POINTER(_f), _vp],
    "avn_clip_resize_transform_u8_bf16": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, C.POINTER(_f), C.POINTER(_f), _vp],
    "avn_layernorm_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _f, _vp],
    "avn_layernorm_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _vp],
    "avn_gemm_bf16": [C.POINTER(GemmDesc), _vp],
    "avn_sizeof_gemm_desc": [],
    "avn_attn_fwd": [_vp, _vp, _vp, _i, _i, _i, _f, _i, _vp],
    "avn_attn_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _vp],
    "avn_cast_f32_bf16": [_vp, _vp, _i64, _vp],
    "avn_transpose_cast_f32_bf16": [_vp, _vp, _i, _i, _vp],
    "avn_colsum_bf16": [_vp, _i64, _vp, _i64, _i, _vp],
    "avn_pos_table_fwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "avn_broadcast_row": [_vp, _vp, _i, _i64, _i, _vp],
    "avn_pos_table_bwd": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "avn_token_embed_fwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "avn_token_embed_bwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "avn_head_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp],
    "avn_head_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "avn_infonce_fwd": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
    "avn_infonce_bwd": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _f, _i, _vp, _vp, _vp, _vp],
    "avn_infonce_static_max_floats": [],
    "avn_infonce_bwd_static": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _f, _i, _vp, _vp, _vp],
    "avn_adamw_chunk_elems": [],
    "avn_adamw_step": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _f, _f, _f, _i, _vp],
    "avn_clamp_scalar_f32": [_vp, _f, _f, _vp],
    "avn_tube_mask_slots": [_vp, _vp, _vp, _vp, _i, _i, _vp],
    "avn_tubelet_transform_u8": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, C.POINTER(_f), C.POINTER(_f), _vp],
    "avn_gather_rows_f32_bf16": [_vp, _vp, _i64, _vp, _i, _vp, _i, _i, _i, _vp],
    "avn_colsum_batched_bf16": [_vp, _i, _i64, _i, _i, _vp, _vp],
    "avn_mse_workspace_floats": [],
    "avn_mse_fwd": [_vp, _vp, _i64, _vp, _vp, _vp],
    "avn_mse_bwd": [_vp, _vp, _vp, _i64, _vp, _vp],
    "avn_mmr_workspace_floats": [_i, _i],
    "avn_mmr_fwd": [_vp, _vp, _i, _i, _f, _i, _vp, _vp, _vp],
    "avn_mmr_bwd": [_vp, _vp, _i, _i, _f, _i, _vp, _vp, _vp],
}

# kernels launched per C-ABI call (for the bench's gpu_launches count); default 1
KERNELS_PER_CALL = {"avn_attn_bwd": 3, "avn_infonce_bwd": 3, "avn_infonce_bwd_static": 3, "avn_infonce_static_max_floats": 0, "avn_head_bwd": 2, "avn_pos_table_bwd": 2,
                    "avn_token_embed_bwd": 2, "avn_sample_rrc_boxes": 0, "avn_device_check": 0, "avn_version": 0, "avn_sizeof_gemm_desc": 0,
                    "avn_adamw_chunk_elems": 0, "avn_mse_workspace_floats": 0, "avn_mse_fwd": 2,
                    "avn_mmr_workspace_floats": 0, "avn_mmr_fwd": 3}
launch_count = 0

_lib = None


def load():
    """Load the library once; raise if it has not been built (python __graft_entry__.py / csrc/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"avion_b200: {LIB_PATH} is missing. Build it with `python avion_b200/csrc/build.py`; "
                "there is no CPU or PyTorch fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = _i
        lib.avn_last_error.restype = C.c_char_p
        lib.avn_last_error.argtypes = []
        if lib.avn_sizeof_gemm_desc() != C.sizeof(GemmDesc):
            raise RuntimeError(f"avion_b200: {LIB_PATH} was built with a {lib.avn_sizeof_gemm_desc()}-byte avn_gemm_desc, this "
                               f"binding mirrors {C.sizeof(GemmDesc)} bytes - rebuild the library (python avion_b200/csrc/build.py)")
        _lib = lib
    return _lib


def call(name, *args):
    global launch_count
    lib = load()
    launch_count += KERNELS_PER_CALL.get(name, 1)
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.avn_last_error().decode()}")
    return rc
