"""ctypes binding of libb200ir.so (the C ABI declared in include/b200ir.h).

There is NO fallback: if the shared library is missing or a symbol is absent the import
fails loudly, and every op raises if it is handed a non-CUDA tensor.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200IR_LIB") or os.path.join(_HERE, "libb200ir.so")   # B200IR_LIB: A/B builds (tools/)

F32, BF16 = 0, 1
LN_NONE, LN_WITHBIAS, LN_BIASFREE = 0, 1, 2
SHUFFLE_NONE, UNSHUFFLE2, SHUFFLE2 = 0, 1, 2
IMPL_AUTO, IMPL_SIMT, IMPL_TC = 0, 1, 2
MDTA_BWD_CHAIN, MDTA_BWD_FOLD = 1, 2
GEMM_AUTO, GEMM_RESIDENT, GEMM_STREAM, GEMM_RESIDENT_CH = 0, 1, 2, 3

_vp, _i64, _i, _sz, _f = C.c_void_p, C.c_int64, C.c_int, C.c_size_t, C.c_float

# name -> argtypes; every symbol of include/b200ir.h must be listed here (tests/test_abi.py checks)
SIGNATURES = {
    "b200ir_version": [],
    "b200ir_last_error_string": [],
    "b200ir_ln_conv1x1_fwd": [_vp, _i64, _i, _vp, _i64, _vp, _i, _vp, _vp, _vp, _i64, _i, _vp, _i64, _i,
                              _i, _i, _i, _i64, _vp],
    "b200ir_dwconv3x3_fwd": [_vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv_kxk_fwd": [_vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv_kxk_bwd_input": [_vp, _i64, _vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv_kxk_wgrad_nparts": [_i, _i, _i],
    "b200ir_dwconv_kxk_wgrad": [_vp, _i64, _vp, _i64, _i, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_sigmoid_gate_fwd": [_vp, _i64, _vp, _i64, _i, _vp, _i64, _i, _i, _i, _i64, _vp],
    "b200ir_sigmoid_gate_bwd": [_vp, _i64, _vp, _i64, _vp, _i64, _i, _vp, _i64, _vp, _i64, _i, _i, _i, _i64, _vp],
    "b200ir_dwconv3x3_bwd_input": [_vp, _i64, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv3x3_gelu_gate_fwd": [_vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp],
    "b200ir_mdta_gram_nslab": [_i, _i, _i, _i64],
    "b200ir_mdta_gram_fwd": [_vp, _i64, _i, _i, _i, _i, _i64, _vp, _i, _i, _vp],
    "b200ir_mdta_softmax_fold_fwd": [_vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "b200ir_mdta_softmax_fold_fwd_ex": [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv3x3_gelu_sigmoid_gate_fwd": [_vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp],
    "b200ir_conv3x3_fwd": [_vp, _i64, _i, _vp, _vp, _vp, _i64, _i, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_copy_cast": [_vp, _i64, _i, _vp, _i64, _i, _i, _i64, _vp],
    "b200ir_gemm_tc_geometry": [_i, _i, C.POINTER(C.c_int), C.POINTER(C.c_int)],
    "b200ir_pack_conv1x1": [_vp, _i64, _i, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _vp],
    "b200ir_pack_conv1x1_t": [_vp, _i, _i, _vp, _vp, _vp],
    "b200ir_pack_conv1x1_multi": [_vp, _i, _i, _vp],
    "b200ir_conv3x3_tc_geometry": [_i, _i, C.POINTER(C.c_int), C.POINTER(C.c_int)],
    "b200ir_nchw_to_nhwc": [_vp, _i64, _vp, _i, _i, _i64, _vp],
    "b200ir_pack_conv3x3": [_vp, _vp, _i, _i, _vp, _vp, _vp],
    "b200ir_conv3x3_tc_fwd": [_vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_gemm_tc_fwd": [_vp, _i64, _vp, _i, _vp, _i, _vp, _i64, _i, _vp, _i64, _i, _i, _i, _i, _i64, _vp],
    "b200ir_gemm_tc_fwd_ex": [_vp, _i64, _vp, _i, _vp, _i, _vp, _i64, _i, _vp, _i64, _i, _i, _i, _i, _i64, _i, _vp],
    "b200ir_fused_proj_dw_geometry": [_i, _i, _i, C.POINTER(C.c_int), C.POINTER(C.c_int)],
    "b200ir_pack_fused_proj_dw": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp],
    "b200ir_ln_proj_dw_supported": [_i, _i, _i, _i, _i],
    "b200ir_ln_proj_dw_fwd": [_vp, _i64, _vp, _vp, _vp, _i, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_ln_fwd": [_vp, _i64, _i, _i, _vp, _vp, _vp, _i64, _i, _i, _i, _i64, _vp],
    "b200ir_ln_bwd_nparts": [_i, _i64],
    "b200ir_ln_bwd": [_vp, _i64, _i, _vp, _i64, _i, _i, _vp, _vp, _vp, _i64, _i, _vp, _i64, _i, _vp, _i64, _vp, _i, _i, _i,
                      _i64, _vp],
    "b200ir_reduce_partials": [_vp, _vp, _i, _i64, _i64, _vp],
    "b200ir_reduce_partials_taps": [_vp, _vp, _vp, _i, _i, _i, _vp],
    "b200ir_conv_wgrad_nsplit": [_i, _i, _i, _i, _i64, _i],
    "b200ir_conv_wgrad": [_vp, _i64, _i, _vp, _i64, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_im2col3x3": [_vp, _i64, _i, _vp, _i, _i, _i, _i, _vp],
    "b200ir_dwconv3x3_bwd_nparts": [_i, _i, _i, _i],
    "b200ir_dwconv3x3_wgrad": [_vp, _i64, _vp, _i64, _i, _vp, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv3x3_gelu_gate_bwd": [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _i, _vp, _i, _i, _i, _i, _i, _vp],
    "b200ir_dwconv3x3_gelu_sigmoid_gate_bwd": [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _i, _vp, _i, _i, _i, _i, _i, _vp],
    "b200ir_mdta_bwd": [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "b200ir_mdta_bwd_ex": [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "b200ir_mdta_bwd_parts": [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "b200ir_pixel_shuffle": [_vp, _i64, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_l1_loss_fwd": [_vp, _i, _vp, _i, _i64, _vp, _i, _vp, _vp],
    "b200ir_l1_loss_fwd_noalloc": [_vp, _i, _vp, _i, _i64, _vp, _vp],
    "b200ir_l1_loss_bwd": [_vp, _i, _vp, _i, _i64, _vp, _vp, _i, _vp],
    "b200ir_circconv_patch": [_vp, _i64, _vp, _i64, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_ew": [_vp, _i64, _vp, _i64, _vp, _i64, _i, _i, _i64, _i, _vp],
    "b200ir_scale_rows": [_vp, _i64, _vp, _vp, _i64, _i, _i, _i64, _i, _vp],
    "b200ir_spatial_mean": [_vp, _i64, _i, _vp, _i, _i, _i64, _i, _vp],
    "b200ir_spatial_mean_bwd": [_vp, _vp, _i64, _vp, _i64, _i, _i, _i, _i64, _i, _vp],
    "b200ir_rows_copy": [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i, _i64, _i, _i, _vp],
    "b200ir_rowdot": [_vp, _i64, _vp, _vp, _i64, _vp, _vp, _i, _i64, _i, _vp],
    "b200ir_lowfreq_split": [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_bilinear_resize": [_vp, _i64, _vp, _i64, _i, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_spatial_gate": [_vp, _i64, _i, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "b200ir_spatial_max": [_vp, _i64, _i, _vp, _i, _i, _i64, _vp],
    "b200ir_fre_combine": [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i64, _vp],
    "b200ir_affine2": [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i64, _vp],
    "b200ir_lowfreq_split_bwd": [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_spatial_gate_bwd": [_vp, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _i, _i, _i, _i, _i, _vp],
    "b200ir_spatial_max_bwd": [_vp, _i64, _vp, _vp, _i64, _i, _i, _i, _i64, _vp],
    "b200ir_fre_combine_bwd": [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i, _i, _i64, _vp],
    "b200ir_affine2_bwd": [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _vp, _i, _i, _i, _i64, _vp],
    "b200ir_crop_augment_degrade": [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
    "b200ir_tile_gather": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_tile_blend": [_vp, _vp, _i, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "b200ir_psnr_ssim_nblocks": [_i, _i],
    "b200ir_psnr_ssim": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "b200ir_adamw_step": [_vp, _vp, _vp, _vp, _i64, _f, _f, _f, _f, _f, _i, _vp, _f, _vp],
    "b200ir_adamw_step_ex": [_vp, _vp, _vp, _vp, _i64, _f, _f, _f, _f, _f, _i, _vp, _i, _vp, _f, _vp],
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python __graft_entry__.py` or "
            f"`image-restoration_b200/csrc/build.sh` (there is no CPU/PyTorch fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: loud by design
        fn.argtypes = argtypes
        fn.restype = C.c_char_p if name == "b200ir_last_error_string" else C.c_int
    return lib


lib = _load()


def check(rc: int, what: str):
    if rc != 0:
        msg = lib.b200ir_last_error_string()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")


class PackDesc(C.Structure):
    """ctypes mirror of b200ir_pack_desc_t (include/b200ir.h)."""
    _fields_ = [("w", C.c_void_p), ("bias", C.c_void_p), ("ln_w", C.c_void_p), ("ln_b", C.c_void_p),
                ("wp", C.c_void_p), ("st", C.c_void_p),
                ("Cin", C.c_int), ("Cout", C.c_int), ("Kpad", C.c_int), ("Nrows", C.c_int),
                ("ln_type", C.c_int), ("transposed", C.c_int), ("first_block", C.c_int),
                ("reserved", C.c_int)]

