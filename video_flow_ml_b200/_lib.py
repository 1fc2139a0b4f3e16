"""ctypes binding of libvfml.so (include/vfml.h).  There is no fallback: if the library is
missing or a call fails, a RuntimeError is raised."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VFML_LIB_PATH") or os.path.join(_HERE, "libvfml.so")     # override: A/B builds of the same ABI

F32, F16 = 0, 1
NCHW, ROWS = 0, 1
EPI_BIAS, EPI_RES, EPI_GELU, EPI_RELU, EPI_DW1, EPI_ADDTO, EPI_WEIGHTS = 1, 2, 4, 8, 16, 32, 64

vp, i32, i64, f32, f64, u32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_uint32, C.c_size_t


class GemmDesc(C.Structure):
    """struct vfml_gemm_desc"""
    _fields_ = [("A", vp), ("B", vp), ("D", vp), ("bias", vp), ("res", vp), ("dw_w", vp), ("dw_b", vp),
                ("lda", i64), ("ldb", i64), ("ldd", i64), ("ldr", i64),
                ("M", i32), ("N", i32), ("K", i32), ("batch", i32),
                ("strideA", i64), ("strideB", i64), ("strideD", i64), ("strideR", i64),
                ("alpha", f32), ("epi", u32), ("in_dtype", i32), ("out_dtype", i32)]


# name -> (restype, argtypes); mirrors include/vfml.h one to one
PROTOTYPES = {
    "vfml_version": (i32, []),
    "vfml_last_error": (C.c_char_p, []),
    "vfml_launch_count": (C.c_uint64, []),
    "vfml_device_sms": (i32, []),
    "vfml_nchw_to_rows": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "vfml_rows_to_nchw": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "vfml_corr_volume_workspace": (sz, [i32, i32, i32, i32, i32]),
    "vfml_corr_volume_pyramid": (i32, [vp, vp, i32, i32, i32, i32, C.POINTER(vp), i32, i32, vp, sz, vp]),
    "vfml_corr_lookup": (i32, [C.POINTER(vp), i32, i32, i32, i32, i32, i32, vp, vp, i32, i32, i32, i32, vp]),
    "vfml_corr_lookup_pair": (i32, [C.POINTER(vp), C.POINTER(vp), i32, i32, i32, i32, i32, vp, vp, vp, vp, i32, i32, i32, vp]),
    "vfml_alt_corr_forward": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "vfml_corr_tiled_slab_bytes": (sz, [i32, i32, i32]),
    "vfml_corr_tile_pyramid": (i32, [vp, i64, i32, i32, C.POINTER(vp), i32, vp]),
    "vfml_corr_volume_tiled_workspace": (sz, [i32, i32, i32, i32]),
    "vfml_corr_volume_pyramid_tiled": (i32, [vp, vp, i32, i32, i32, i32, C.POINTER(vp), i32, vp, sz, vp]),
    "vfml_corr_volume_pyramid_tiled_static_max_bytes": (sz, []),
    "vfml_corr_volume_pyramid_tiled_static": (i32, [vp, vp, i32, i32, i32, i32, C.POINTER(vp), i32, vp]),
    "vfml_corr_lookup_tiled": (i32, [C.POINTER(vp), C.POINTER(vp), i32, i32, i32, vp, vp, vp, vp, i32, i32, i32, vp]),
    "vfml_convex_upsample8": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, i64, i64, i64, i32, i32, i32, i32, i32, vp]),
    "vfml_encode_gamedev": (i32, [vp, vp, i32, i32, f32, f32, f32, f32, vp]),
    "vfml_encode_mv_rg8": (i32, [vp, vp, i32, i32, f32, vp]),
    "vfml_decode_mv_rg8": (i32, [vp, vp, i32, i32, f32, vp]),
    "vfml_encode_mv_rgb8": (i32, [vp, vp, i32, i32, f32, vp]),
    "vfml_decode_mv_rgb8": (i32, [vp, vp, i32, i32, f32, vp]),
    "vfml_encode_hsv": (i32, [vp, vp, i32, i32, vp, vp]),
    "vfml_encode_torchvision": (i32, [vp, vp, i32, i32, i32, vp, vp]),
    "vfml_lod_pyramid": (i32, [vp, i32, i32, C.POINTER(vp), i32, vp]),
    "vfml_quality_frame": (i32, [vp, vp, vp, i32, i32, vp, i32, i32, f32, vp]),
    "vfml_taa": (i32, [vp, i32, vp, vp, vp, i32, i32, f64, f64, i32, i32, vp]),
    "vfml_taa_init": (i32, [vp, i32, vp, i32, i32, vp]),
    "vfml_gemm": (i32, [C.POINTER(GemmDesc), vp]),
    "vfml_gemm_autotune": (i32, [C.POINTER(GemmDesc), vp]),
    "vfml_gemm_split_weight_bytes": (sz, [i32, i32]),
    "vfml_gemm_split_weight": (i32, [vp, i32, i32, i64, vp, vp]),
    "vfml_gemm_split_workspace": (sz, [i32, i32, i32]),
    "vfml_gemm_f32_split": (i32, [C.POINTER(GemmDesc), vp, vp, sz, vp]),
    "vfml_gemm_uses_tensor_cores": (i32, [C.POINTER(GemmDesc)]),
    "vfml_dwconv_res_gelu": (i32, [vp, vp, i32, i32, i32, i32, i64, i64, vp, vp, i32, i32, vp]),
    "vfml_dwconv_packed_bytes": (sz, [i32, i32]),
    "vfml_dwconv_pack_weights": (i32, [vp, i32, i32, vp, vp]),
    "vfml_dwconv_tc_res_gelu": (i32, [vp, vp, i32, i32, i32, i32, i64, i64, vp, vp, i32, vp]),
    "vfml_splitk_reduce": (i32, [vp, i32, i64, i32, vp, i64, vp, i64, i32, vp]),
    "vfml_layer_norm_rows": (i32, [vp, vp, i64, i32, i64, i64, vp, vp, f32, i32, vp]),
    "vfml_attention_hd32": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i64, i64, i64, i64, i64, i64, f32, vp]),
    "vfml_peg3x3": (i32, [vp, vp, i32, i32, i32, i32, i64, i64, vp, vp, i32, vp]),
    "vfml_softmax_rows": (i32, [vp, vp, i64, i32, i64, i64, i32, vp]),
    "vfml_motion_glue": (i32, [vp, vp, vp, i64, vp, i64, vp, i64, i32, vp, vp, i32, i32, i32, i32, i32, vp]),
    "vfml_coords_add": (i32, [vp, vp, vp, i64, i32, i32, vp]),
    "vfml_coords_grid": (i32, [vp, i32, i32, i32, vp]),
    "vfml_rows_fill": (i32, [vp, i64, i32, i32, i64, vp, i32, vp]),
    "vfml_im2col3x3": (i32, [vp, vp, i32, i32, i32, i32, i64, i64, i32, vp]),
}

_lib = None


def load():
    """Load libvfml.so once and attach prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m video_flow_ml_b200.build` "
            "(there is no CPU or PyTorch fallback for the CUDA hot path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)          # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.vfml_version() != 100:
        raise RuntimeError(f"libvfml version {lib.vfml_version()} does not match the binding (100)")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().vfml_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libvfml {what} failed ({rc}): {msg}")


def launch_count():
    return int(load().vfml_launch_count())
