"""Tensor-level wrappers over the C ABI (include/vfml.h).  torch is used for device memory and
streams only; every function launches libvfml kernels on the current CUDA stream.

"rows views": 2-D tensors [n_rows, n_cols] with stride(1) == 1 (column slices of a wider
[n_rows, pitch] buffer are fine); their row stride is the pitch handed to the kernels.
"""
import ctypes as C
import functools

import os

import torch

from . import _lib
from ._lib import F16, F32, NCHW, ROWS

_DT = {torch.float32: F32, torch.float16: F16}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("libvfml operates on CUDA tensors only (no CPU fallback)")


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _dt(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise ValueError(f"unsupported dtype {t.dtype}") from None


def _rows(t, name):
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise ValueError(f"{name}: expected a rows view [n, c] with unit column stride, got {tuple(t.shape)} / {t.stride()}")
    return t


def _f32c(t, name):
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise ValueError(f"{name}: expected a contiguous float32 tensor")
    return t


# ------------------------------------------------------------------ layout
def nchw_to_rows(src, dst):
    """src [n, c, h, w] fp32 -> dst rows view [n*h*w, c]."""
    _cuda(src, dst)
    _f32c(src, "src"); _rows(dst, "dst")
    n, c, h, w = src.shape
    if dst.shape != (n * h * w, c):
        raise ValueError("nchw_to_rows: shape mismatch")
    _lib.check(_lib.load().vfml_nchw_to_rows(_p(src), _p(dst), n, c, h * w, dst.stride(0), 0, _dt(dst), _stream()), "nchw_to_rows")
    return dst


def rows_to_nchw(src, dst):
    """src rows view [n*h*w, c] -> dst [n, c, h, w] fp32."""
    _cuda(src, dst)
    _rows(src, "src"); _f32c(dst, "dst")
    n, c, h, w = dst.shape
    if src.shape != (n * h * w, c):
        raise ValueError("rows_to_nchw: shape mismatch")
    _lib.check(_lib.load().vfml_rows_to_nchw(_p(src), _p(dst), n, c, h * w, src.stride(0), 0, _dt(src), _stream()), "rows_to_nchw")
    return dst


# ------------------------------------------------------------------ correlation
def pyramid_shapes(h, w, num_levels):
    out = []
    for _ in range(num_levels):
        out.append((h, w))
        h, w = h // 2, w // 2
    return out


def corr_volume_pyramid(fmap1, fmap2, num_levels=4, vol_dtype=torch.float32):
    """-> list of level tensors [n*h*w, 1, h_l, w_l] (the reference's CorrBlock.corr_pyramid)."""
    _cuda(fmap1, fmap2)
    _f32c(fmap1, "fmap1"); _f32c(fmap2, "fmap2")
    if fmap1.shape != fmap2.shape or fmap1.dim() != 4:
        raise ValueError("corr_volume_pyramid: fmaps must be [n, C, h, w] of equal shape")
    n, Cc, h, w = fmap1.shape
    levels = [torch.empty((n * h * w, 1, hl, wl), dtype=vol_dtype, device=fmap1.device)
              for hl, wl in pyramid_shapes(h, w, num_levels)]
    if any(l.numel() == 0 for l in levels) and n > 0:
        raise ValueError(f"corr_volume_pyramid: {h}x{w} is too small for {num_levels} levels")
    arr = (C.c_void_p * num_levels)(*[l.data_ptr() for l in levels])
    lib = _lib.load()
    ws_bytes = lib.vfml_corr_volume_workspace(n, Cc, h, w, _DT[vol_dtype])
    ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=fmap1.device)
    _lib.check(lib.vfml_corr_volume_pyramid(_p(fmap1), _p(fmap2), n, Cc, h, w, arr, num_levels, _DT[vol_dtype],
                                            _p(ws), ws_bytes, _stream()), "corr_volume_pyramid")
    return levels


def corr_lookup(levels, coords, radius=4, out=None, out_dtype=torch.float32):
    """coords [n, 2, h, w] fp32.  out=None -> new NCHW [n, L*(2r+1)^2, h, w];
    out = rows view [n*h*w, L*(2r+1)^2] -> written in rows layout."""
    _cuda(coords, *levels)
    _f32c(coords, "coords")
    n, two, h, w = coords.shape
    L = len(levels)
    ch = L * (2 * radius + 1) ** 2
    if two != 2 or levels[0].shape[0] != n * h * w or tuple(levels[0].shape[-2:]) != (h, w):
        raise ValueError("corr_lookup: coords do not match the pyramid")
    arr = (C.c_void_p * L)(*[l.data_ptr() for l in levels])
    if out is None:
        out = torch.empty((n, ch, h, w), dtype=out_dtype, device=coords.device)
        layout, pitch = NCHW, 0
    else:
        _rows(out, "out")
        if out.shape != (n * h * w, ch):
            raise ValueError("corr_lookup: out must be [n*h*w, L*(2r+1)^2]")
        layout, pitch = ROWS, out.stride(0)
    _lib.check(_lib.load().vfml_corr_lookup(arr, L, _dt(levels[0]), n, h, w, radius, _p(coords), _p(out), layout, pitch, 0,
                                            _dt(out), _stream()), "corr_lookup")
    return out


def alt_corr_forward(fmap1, fmap2, coords, radius):
    """alt_cuda_corr.forward (correlation.cpp:23-33): fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C] NHWC contiguous (fp32 or fp16),
    coords [B,N,H1,W1,2] fp32 -> [corr [B,N,(2r+1)^2,H1,W1] fp32] (a one-element list, like the reference binding)."""
    _cuda(fmap1, fmap2, coords)
    _f32c(coords, "coords")
    if fmap1.dim() != 4 or fmap2.dim() != 4 or coords.dim() != 5 or coords.shape[-1] != 2:
        raise ValueError("alt_corr_forward: expected fmap [B,H,W,C] and coords [B,N,H1,W1,2]")
    for t, name in ((fmap1, "fmap1"), (fmap2, "fmap2")):
        if not t.is_contiguous():                      # CHECK_CONTIGUOUS, correlation.cpp:19-21
            raise ValueError(f"alt_corr_forward: {name} must be contiguous")
    B, H1, W1, Cc = fmap1.shape
    if fmap2.shape[0] != B or fmap2.shape[3] != Cc or fmap2.dtype != fmap1.dtype or tuple(coords.shape[2:4]) != (H1, W1) or coords.shape[0] != B:
        raise ValueError("alt_corr_forward: shape mismatch")
    N = coords.shape[1]
    corr = torch.empty((B, N, (2 * radius + 1) ** 2, H1, W1), dtype=torch.float32, device=fmap1.device)
    _lib.check(_lib.load().vfml_alt_corr_forward(_p(fmap1), _p(fmap2), _p(coords), _p(corr), B, N, H1, W1, fmap2.shape[1],
                                                 fmap2.shape[2], Cc, radius, _dt(fmap1), _stream()), "alt_corr_forward")
    return [corr]


def corr_lookup_pair(levels_f, levels_b, coords_f, coords_b, out_f, out_b):
    """Forward + backward radius-4 lookups in one launch; out_f / out_b: rows views [n*h*w, 324] of equal pitch."""
    _cuda(coords_f, coords_b, out_f, out_b, *levels_f, *levels_b)
    _f32c(coords_f, "coords_f"); _f32c(coords_b, "coords_b")
    _rows(out_f, "out_f"); _rows(out_b, "out_b")
    n, _, h, w = coords_f.shape
    L = len(levels_f)
    if (coords_b.shape != coords_f.shape or len(levels_b) != L or out_f.shape != (n * h * w, L * 81)
            or out_b.shape != out_f.shape or out_f.stride(0) != out_b.stride(0) or out_f.dtype != out_b.dtype
            or levels_f[0].shape != levels_b[0].shape or levels_f[0].shape[0] != n * h * w):
        raise ValueError("corr_lookup_pair: shape mismatch")
    af = (C.c_void_p * L)(*[l.data_ptr() for l in levels_f])
    ab = (C.c_void_p * L)(*[l.data_ptr() for l in levels_b])
    _lib.check(_lib.load().vfml_corr_lookup_pair(af, ab, L, _dt(levels_f[0]), n, h, w, _p(coords_f), _p(coords_b), _p(out_f),
                                                 _p(out_b), out_f.stride(0), 0, _dt(out_f), _stream()), "corr_lookup_pair")


class TiledPyramid:
    """The engine's gather-friendly fp16 pyramid (include/vfml.h, "tiled" layout): per level a uint8
    buffer of [n*h*w] slabs of 4x4-texel (32-byte) tiles."""

    def __init__(self, levels, n, h, w):
        self.levels, self.n, self.h, self.w = levels, n, h, w


def corr_lookup_perm(device=None):
    """perm[new] = old: column `new` of the tiled lookup holds the reference's channel `old`
    (l*81 + i*9 + j); new = (l*9+i)*8 + j for j < 8, 288 + l*9 + i for j == 8."""
    perm = torch.empty(324, dtype=torch.long)
    for l in range(4):
        for i in range(9):
            for j in range(9):
                new = (l * 9 + i) * 8 + j if j < 8 else 288 + l * 9 + i
                perm[new] = l * 81 + i * 9 + j
    return perm.to(device) if device is not None else perm


def _tiled_levels(n_rows, h, w, device):
    lib = _lib.load()
    return [torch.empty((n_rows, lib.vfml_corr_tiled_slab_bytes(h, w, l)), dtype=torch.uint8, device=device) for l in range(4)]


def corr_tile_pyramid(vol0, h, w):
    """vol0 [rows, h*w] fp16 (row-major level-0 volume) -> TiledPyramid with `rows` source pixels."""
    _cuda(vol0)
    if vol0.dtype != torch.float16 or not vol0.is_contiguous() or vol0.dim() != 2 or vol0.shape[1] != h * w:
        raise ValueError("corr_tile_pyramid: vol0 must be a contiguous fp16 [rows, h*w] tensor")
    if h < 8 or w < 8:
        raise ValueError(f"corr_tile_pyramid: {h}x{w} is too small for 4 levels")
    rows = vol0.shape[0]
    levels = _tiled_levels(rows, h, w, vol0.device)
    arr = (C.c_void_p * 4)(*[l.data_ptr() for l in levels])
    _lib.check(_lib.load().vfml_corr_tile_pyramid(_p(vol0), rows, h, w, arr, 4, _stream()), "corr_tile_pyramid")
    return TiledPyramid(levels, rows // (h * w), h, w)


def tiled_pyramid_alloc(n, h, w, device):
    """Uninitialised TiledPyramid for n frames (to be filled by corr_volume_pyramid_tiled(..., out=))."""
    return TiledPyramid(_tiled_levels(n * h * w, h, w, device), n, h, w)


def tiled_volume_workspace(n, Cc, h, w, device):
    return torch.empty(max(_lib.load().vfml_corr_volume_tiled_workspace(n, Cc, h, w), 1), dtype=torch.uint8, device=device)


def corr_volume_pyramid_tiled(fmap1, fmap2, workspace=None, out=None):
    """fmaps [n, C, h, w] fp32 -> TiledPyramid (tcgen05 GEMM + fused pool/tile pass); `out`: a TiledPyramid to fill."""
    _cuda(fmap1, fmap2)
    _f32c(fmap1, "fmap1"); _f32c(fmap2, "fmap2")
    if fmap1.shape != fmap2.shape or fmap1.dim() != 4:
        raise ValueError("corr_volume_pyramid_tiled: fmaps must be [n, C, h, w] of equal shape")
    n, Cc, h, w = fmap1.shape
    if h < 8 or w < 8:
        raise ValueError(f"corr_volume_pyramid_tiled: {h}x{w} is too small for 4 levels")
    lib = _lib.load()
    if out is not None and (out.n, out.h, out.w) != (n, h, w):
        raise ValueError("corr_volume_pyramid_tiled: `out` was allocated for another shape")
    levels = out.levels if out is not None else _tiled_levels(n * h * w, h, w, fmap1.device)
    arr = (C.c_void_p * 4)(*[l.data_ptr() for l in levels])
    ws_bytes = lib.vfml_corr_volume_tiled_workspace(n, Cc, h, w)
    if workspace is None or workspace.numel() < ws_bytes:
        workspace = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=fmap1.device)
    _lib.check(lib.vfml_corr_volume_pyramid_tiled(_p(fmap1), _p(fmap2), n, Cc, h, w, arr, 4, _p(workspace), workspace.numel(),
                                                  _stream()), "corr_volume_pyramid_tiled")
    return out if out is not None else TiledPyramid(levels, n, h, w)


def corr_lookup_tiled(pyr_a, pyr_b, coords_a, coords_b, out_a, out_b, math_f32=False):
    """Radius-4 lookups on tiled pyramids, one launch for both sets (pyr_b may be None).
    out_*: fp16 rows views [n*h*w, 324] in the PERMUTED channel order (corr_lookup_perm)."""
    _cuda(coords_a, out_a, *pyr_a.levels)
    _f32c(coords_a, "coords_a"); _rows(out_a, "out_a")
    n, _, h, w = coords_a.shape
    if (pyr_a.n, pyr_a.h, pyr_a.w) != (n, h, w) or out_a.shape != (n * h * w, 324) or out_a.dtype != torch.float16:
        raise ValueError("corr_lookup_tiled: shape mismatch")
    aa = (C.c_void_p * 4)(*[l.data_ptr() for l in pyr_a.levels])
    ab = None
    if pyr_b is not None:
        _cuda(coords_b, out_b, *pyr_b.levels)
        _f32c(coords_b, "coords_b"); _rows(out_b, "out_b")
        if ((pyr_b.n, pyr_b.h, pyr_b.w) != (n, h, w) or coords_b.shape != coords_a.shape or out_b.shape != out_a.shape
                or out_b.stride(0) != out_a.stride(0) or out_b.dtype != torch.float16):
            raise ValueError("corr_lookup_tiled: second set does not match the first")
        ab = (C.c_void_p * 4)(*[l.data_ptr() for l in pyr_b.levels])
    _lib.check(_lib.load().vfml_corr_lookup_tiled(aa, ab, n, h, w, _p(coords_a), _p(coords_b), _p(out_a), _p(out_b),
                                                  out_a.stride(0), 0, int(math_f32), _stream()), "corr_lookup_tiled")
    return out_a, out_b


def convex_upsample8(flow, mask, out=None, crop=None, flow_is_coords=False):
    """flow [n, 2, h, w] fp32; mask [n, 576, h, w] fp32 or rows view [n*h*w, 576].
    crop=(top, left, out_h, out_w) selects a window of the 8h x 8w result (unpad fused);
    out: fp32 tensor view [n, 2, out_h, out_w] with unit x stride (may be a slice of a larger frame)."""
    _cuda(flow, mask)
    _f32c(flow, "flow")
    n, _, h, w = flow.shape
    top, left, oh, ow = crop if crop is not None else (0, 0, 8 * h, 8 * w)
    if out is None:
        out = torch.empty((n, 2, oh, ow), dtype=torch.float32, device=flow.device)
    if out.dtype != torch.float32 or out.shape != (n, 2, oh, ow) or (ow > 1 and out.stride(3) != 1):
        raise ValueError("convex_upsample8: bad out tensor")
    if mask.dim() == 4:
        _f32c(mask, "mask")
        if mask.shape != (n, 576, h, w):
            raise ValueError("convex_upsample8: mask must be [n, 576, h, w]")
        layout, pitch = NCHW, 0
    else:
        _rows(mask, "mask")
        if mask.shape != (n * h * w, 576):
            raise ValueError("convex_upsample8: mask rows must be [n*h*w, 576]")
        layout, pitch = ROWS, mask.stride(0)
    _lib.check(_lib.load().vfml_convex_upsample8(_p(flow), _p(mask), layout, pitch, 0, _dt(mask), n, h, w, _p(out),
                                                 out.stride(0), out.stride(1), out.stride(2), top, left, oh, ow,
                                                 int(flow_is_coords), _stream()), "convex_upsample8")
    return out


# ------------------------------------------------------------------ post-pass
def encode_gamedev(flow, width, height, scale=200.0, clamp=20.0):
    """flow [H, W, 2] fp32 CUDA -> [H, W, 3] u8 CUDA."""
    _cuda(flow)
    _f32c(flow, "flow")
    H, W, _ = flow.shape
    out = torch.empty((H, W, 3), dtype=torch.uint8, device=flow.device)
    _lib.check(_lib.load().vfml_encode_gamedev(_p(flow), _p(out), H, W, float(width), float(height), float(scale),
                                               float(clamp), _stream()), "encode_gamedev")
    return out


def encode_mv_rg8(flow, clamp=64.0):
    _cuda(flow)
    _f32c(flow, "flow")
    H, W, _ = flow.shape
    out = torch.empty((H, W, 3), dtype=torch.uint8, device=flow.device)
    _lib.check(_lib.load().vfml_encode_mv_rg8(_p(flow), _p(out), H, W, float(clamp), _stream()), "encode_mv_rg8")
    return out


def decode_mv_rg8(rgb, clamp=64.0):
    _cuda(rgb)
    if rgb.dtype != torch.uint8 or not rgb.is_contiguous():
        raise ValueError("decode_mv_rg8: expected contiguous uint8")
    H, W, _ = rgb.shape
    out = torch.empty((H, W, 2), dtype=torch.float32, device=rgb.device)
    _lib.check(_lib.load().vfml_decode_mv_rg8(_p(rgb), _p(out), H, W, float(clamp), _stream()), "decode_mv_rg8")
    return out


def _u8_like(flow):
    H, W, two = flow.shape
    if two != 2:
        raise ValueError("flow must be [H, W, 2]")
    return H, W, torch.empty((H, W, 3), dtype=torch.uint8, device=flow.device)


def encode_mv_rgb8(flow, clamp=32.0):
    """flow [H, W, 2] fp32 CUDA -> [H, W, 3] u8: direction (R, G) + sqrt(1 - |d|^2) (B), flow_encoders.py:242-262."""
    _cuda(flow)
    _f32c(flow, "flow")
    H, W, out = _u8_like(flow)
    _lib.check(_lib.load().vfml_encode_mv_rgb8(_p(flow), _p(out), H, W, float(clamp), _stream()), "encode_mv_rgb8")
    return out


def decode_mv_rgb8(rgb, clamp=32.0):
    _cuda(rgb)
    if rgb.dtype != torch.uint8 or not rgb.is_contiguous() or rgb.dim() != 3 or rgb.shape[2] != 3:
        raise ValueError("decode_mv_rgb8: expected contiguous uint8 [H, W, 3]")
    H, W, _ = rgb.shape
    out = torch.empty((H, W, 2), dtype=torch.float32, device=rgb.device)
    _lib.check(_lib.load().vfml_decode_mv_rgb8(_p(rgb), _p(out), H, W, float(clamp), _stream()), "decode_mv_rgb8")
    return out


def encode_hsv(flow):
    """HSVFlowEncoder.encode (flow_encoders.py:38-67) incl. OpenCV's HSV -> RGB."""
    _cuda(flow)
    _f32c(flow, "flow")
    H, W, out = _u8_like(flow)
    scratch = torch.empty(4, dtype=torch.uint8, device=flow.device)
    _lib.check(_lib.load().vfml_encode_hsv(_p(flow), _p(out), H, W, _p(scratch), _stream()), "encode_hsv")
    return out


def encode_torchvision(flow, ref_wrap=True):
    """TorchvisionFlowEncoder.encode (flow_encoders.py:399-427).  ref_wrap=False gives torchvision's flow_to_image itself."""
    _cuda(flow)
    _f32c(flow, "flow")
    H, W, out = _u8_like(flow)
    scratch = torch.empty(4, dtype=torch.uint8, device=flow.device)
    _lib.check(_lib.load().vfml_encode_torchvision(_p(flow), _p(out), H, W, int(bool(ref_wrap)), _p(scratch), _stream()),
               "encode_torchvision")
    return out


def lod_pyramid(flow, num_lods=5):
    """LODGenerator.generate_lods (storage/cache_manager.py:77-161): [flow, lod1, ...], level l+1 = [ceil(h/2), ceil(w/2), 2]."""
    _cuda(flow)
    _f32c(flow, "flow")
    if flow.dim() != 3 or flow.shape[2] != 2 or flow.shape[0] < 1 or flow.shape[1] < 1:
        raise ValueError("lod_pyramid: flow must be a non-empty [H, W, 2] tensor")
    lods = [flow]
    while len(lods) < num_lods:
        src = lods[-1]
        n = min(4, num_lods - len(lods))
        h, w = src.shape[:2]
        new = []
        for _ in range(n):
            h, w = (h + 1) // 2, (w + 1) // 2
            new.append(torch.empty((h, w, 2), dtype=torch.float32, device=flow.device))
        arr = (C.c_void_p * n)(*[t.data_ptr() for t in new])
        _lib.check(_lib.load().vfml_lod_pyramid(_p(src), src.shape[0], src.shape[1], arr, n, _stream()), "lod_pyramid")
        lods += new
    return lods


def quality_frame(frame1, frame2, flow, threshold):
    """generate_quality_frame_gpu (correction_worker.py:175-208): u8 frames [H, W, 3], fp32 flow [h, w, 2] -> u8 [H, W, 3]."""
    _cuda(frame1, frame2, flow)
    _f32c(flow, "flow")
    for f in (frame1, frame2):
        if f.dtype != torch.uint8 or not f.is_contiguous() or f.dim() != 3 or f.shape[2] != 3:
            raise ValueError("quality_frame: frames must be contiguous uint8 [H, W, 3]")
    if frame1.shape != frame2.shape or flow.dim() != 3 or flow.shape[2] != 2:
        raise ValueError("quality_frame: shape mismatch")
    H, W, _ = frame1.shape
    out = torch.empty((H, W, 3), dtype=torch.uint8, device=frame1.device)
    _lib.check(_lib.load().vfml_quality_frame(_p(frame1), _p(frame2), _p(flow), flow.shape[0], flow.shape[1], _p(out), H, W,
                                              float(threshold), _stream()), "quality_frame")
    return out


def _taa_cur(cur):
    if cur.dtype not in (torch.uint8, torch.float32) or cur.dim() != 3 or cur.shape[2] != 3:
        raise ValueError("taa: the current frame must be a uint8 or float32 [H, W, 3] tensor")
    return cur.contiguous(), int(cur.dtype == torch.float32)


def taa_init(cur):
    _cuda(cur)
    cur, is32 = _taa_cur(cur)
    H, W, _ = cur.shape
    out = torch.empty((H, W, 3), dtype=torch.float64, device=cur.device)
    _lib.check(_lib.load().vfml_taa_init(_p(cur), is32, _p(out), H, W, _stream()), "taa_init")
    return out


def taa(cur, flow, history, alpha, sigma_color, mode, hist_f32):
    """cur [H,W,3] u8 or fp32, flow [H,W,2] fp32 or None, history [H,W,3] f64 -> new [H,W,3] f64."""
    _cuda(cur, flow, history)
    cur, is32 = _taa_cur(cur)
    H, W, _ = cur.shape
    if history.dtype != torch.float64 or history.shape != cur.shape:
        raise ValueError("taa: bad tensor types")
    if flow is not None:
        _f32c(flow, "flow")
    out = torch.empty_like(history)
    _lib.check(_lib.load().vfml_taa(_p(cur), is32, _p(flow), _p(history.contiguous()), _p(out), H, W, float(alpha),
                                    float(sigma_color), int(mode), int(bool(hist_f32)), _stream()), "taa")
    return out


# ------------------------------------------------------------------ update-step building blocks
def gemm(a, w, out, bias=None, res=None, act=None, dw=None, alpha=1.0, addto=False, require_tc=False, weights=False):
    """out[m, n] = epi(alpha * sum_k a[m, k] w[n, k]).  a, w, out, res: rows views."""
    _cuda(a, w, out, bias, res)
    _rows(a, "a"); _rows(w, "w"); _rows(out, "out")
    M, K = a.shape
    N = w.shape[0]
    if w.shape[1] != K or out.shape != (M, N) or a.dtype != w.dtype:
        raise ValueError(f"gemm: shape/dtype mismatch a{tuple(a.shape)} w{tuple(w.shape)} out{tuple(out.shape)}")
    d = _lib.GemmDesc()
    d.A, d.B, d.D = a.data_ptr(), w.data_ptr(), out.data_ptr()
    d.lda, d.ldb, d.ldd = a.stride(0), w.stride(0), out.stride(0)
    d.M, d.N, d.K, d.batch = M, N, K, 1
    d.alpha = alpha
    epi = 0
    if bias is not None:
        _f32c(bias, "bias")
        d.bias = bias.data_ptr(); epi |= _lib.EPI_BIAS
    if res is not None:
        _rows(res, "res")
        if res.shape != (M, N) or res.dtype != a.dtype:
            raise ValueError("gemm: residual mismatch")
        d.res = res.data_ptr(); d.ldr = res.stride(0); epi |= _lib.EPI_RES
    if act == "gelu":
        epi |= _lib.EPI_GELU
    elif act == "relu":
        epi |= _lib.EPI_RELU
    elif act is not None:
        raise ValueError(f"gemm: unknown activation {act}")
    if dw is not None:
        d.dw_w, d.dw_b = _f32c(dw[0], "dw_w").data_ptr(), _f32c(dw[1], "dw_b").data_ptr()
        epi |= _lib.EPI_DW1
    if addto:
        epi |= _lib.EPI_ADDTO
    if weights:           # w, bias, dw are layer constants: fetched while the preceding kernel drains
        epi |= _lib.EPI_WEIGHTS
    d.epi = epi
    d.in_dtype, d.out_dtype = _dt(a), _dt(out)
    if require_tc and not _lib.load().vfml_gemm_uses_tensor_cores(C.byref(d)):
        raise RuntimeError("gemm: descriptor is not eligible for the tensor-core kernel (alignment / dtype)")
    if _GEMM_AUTOTUNE[0] and a.dtype == torch.float16:
        rc = _lib.load().vfml_gemm_autotune(C.byref(d), _stream())
        if rc < 0:
            _lib.check(rc, "gemm_autotune")
    if a.dtype == torch.float32 and out.dtype == torch.float32 and _FP32_SPLIT[0] and not addto and M >= 512 and N >= 16 and K >= 16:
        # the exact tier's products on the tensor cores: fp32 = fp16 hi + fp16 lo, three products in one fp16 GEMM (gemm_split.cu)
        lib = _lib.load()
        sid = torch.cuda.current_stream().cuda_stream      # concurrent branches run on their own streams: scratch per stream
        dev_i = a.device.index
        nb = lib.vfml_gemm_split_weight_bytes(N, K)
        if weights:                    # layer constants: split once, keyed by their storage (the cache keeps the tensor alive)
            wkey = (w.data_ptr(), N, K, w.stride(0), dev_i)
            ent = _SPLIT_WEIGHTS.get(wkey)
            if ent is None:
                w3 = torch.empty(nb, dtype=torch.uint8, device=w.device)
                _lib.check(lib.vfml_gemm_split_weight(_p(w), N, K, w.stride(0), _p(w3), _stream()), "gemm_split_weight")
                while len(_SPLIT_WEIGHTS) >= 512:
                    _SPLIT_WEIGHTS.pop(next(iter(_SPLIT_WEIGHTS)))
                ent = _SPLIT_WEIGHTS[wkey] = (w3, w)
            w3 = ent[0]
        else:                          # an activation as the second operand (q k^T, attn v): split per call
            bkey = ("b", sid, N, K, dev_i)
            w3 = _SPLIT_SCRATCH.get(bkey)
            if w3 is None:
                w3 = _SPLIT_SCRATCH[bkey] = torch.empty(nb, dtype=torch.uint8, device=w.device)
            _lib.check(lib.vfml_gemm_split_weight(_p(w), N, K, w.stride(0), _p(w3), _stream()), "gemm_split_weight")
        skey = ("a", sid, M, N, K, dev_i)
        wsb = _SPLIT_SCRATCH.get(skey)
        if wsb is None:
            wsb = _SPLIT_SCRATCH[skey] = torch.empty(lib.vfml_gemm_split_workspace(M, N, K), dtype=torch.uint8, device=a.device)
        _lib.check(lib.vfml_gemm_f32_split(C.byref(d), _p(w3), _p(wsb), wsb.numel(), _stream()), "gemm_f32_split")
        return out
    _lib.check(_lib.load().vfml_gemm(C.byref(d), _stream()), "gemm")
    return out


# Tensor-core route of fp32 products (opt-in: VFML_FP32_SPLIT=1 or ops.set_fp32_split(True)).  22-bit operands instead of 24:
# 1.2e-4 px from the reference on the well-conditioned fixtures (FFMA: 3.6e-5) at 1.5x the speed of the FFMA tier, but 6e-2 px on
# the ill-conditioned gain-1.4 set where the FFMA kernels hold 2e-3 -- so the default fp32 tier stays exact.
_FP32_SPLIT = [os.environ.get("VFML_FP32_SPLIT", "0") == "1"]


def set_fp32_split(enabled):
    _FP32_SPLIT[0] = bool(enabled)

_SPLIT_WEIGHTS = {}
_SPLIT_SCRATCH = {}


_GEMM_AUTOTUNE = [False]


class gemm_autotuning:
    """with ops.gemm_autotuning(): ...   every ops.gemm inside first lets the library time its kernels for that layer shape
    (vfml_gemm_autotune: synchronises, so only outside graph capture -- the engines use it around their eager warm-up)."""

    def __enter__(self):
        self.prev = _GEMM_AUTOTUNE[0]
        _GEMM_AUTOTUNE[0] = os.environ.get("VFML_GEMM_NO_AUTOTUNE") is None
        return self

    def __exit__(self, *exc):
        _GEMM_AUTOTUNE[0] = self.prev
        return False


def gemm_batched(a, b, out, alpha=1.0):
    """out[i] = alpha * a[i] @ b[i]^T for 3-D views [batch, rows, k] with unit k stride."""
    _cuda(a, b, out)
    Bn, M, K = a.shape
    N = b.shape[1]
    if b.shape != (Bn, N, K) or out.shape != (Bn, M, N) or a.stride(2) != 1 or b.stride(2) != 1 or out.stride(2) != 1:
        raise ValueError("gemm_batched: shape mismatch")
    d = _lib.GemmDesc()
    d.A, d.B, d.D = a.data_ptr(), b.data_ptr(), out.data_ptr()
    d.lda, d.ldb, d.ldd = a.stride(1), b.stride(1), out.stride(1)
    d.strideA, d.strideB, d.strideD = a.stride(0), b.stride(0), out.stride(0)
    d.M, d.N, d.K, d.batch = M, N, K, Bn
    d.alpha = alpha
    d.in_dtype, d.out_dtype = _dt(a), _dt(out)
    _lib.check(_lib.load().vfml_gemm(C.byref(d), _stream()), "gemm_batched")
    return out


def dwconv_res_gelu(x, y, n_img, h, w, weight_tapmajor, bias, k):
    """y = gelu(x + dwconv_k(x) + b) over rows views [n_img*h*w, C]; weight [k*k, C] fp32."""
    _cuda(x, y, weight_tapmajor, bias)
    _rows(x, "x"); _rows(y, "y")
    Cc = x.shape[1]
    if x.shape != y.shape or x.shape[0] != n_img * h * w or weight_tapmajor.shape != (k * k, Cc) or x.dtype != y.dtype:
        raise ValueError("dwconv_res_gelu: shape mismatch")
    _lib.check(_lib.load().vfml_dwconv_res_gelu(_p(x), _p(y), n_img, h, w, Cc, x.stride(0), y.stride(0),
                                                _p(_f32c(weight_tapmajor, "w")), _p(_f32c(bias, "b")), k, _dt(x), _stream()),
               "dwconv_res_gelu")
    return y


def dwconv_pack_weights(weight_tapmajor, k):
    """fp32 tap-major [k*k, C] depthwise weight -> packed fp16 band table for dwconv_tc_res_gelu (uint8 tensor)."""
    _cuda(weight_tapmajor)
    _f32c(weight_tapmajor, "weight")
    if weight_tapmajor.dim() != 2 or weight_tapmajor.shape[0] != k * k:
        raise ValueError("dwconv_pack_weights: weight must be [k*k, C]")
    Cc = weight_tapmajor.shape[1]
    lib = _lib.load()
    nbytes = lib.vfml_dwconv_packed_bytes(Cc, k)
    if nbytes == 0:
        raise ValueError(f"dwconv_pack_weights: unsupported k={k}")
    packed = torch.empty(nbytes, dtype=torch.uint8, device=weight_tapmajor.device)
    _lib.check(lib.vfml_dwconv_pack_weights(_p(weight_tapmajor), Cc, k, _p(packed), _stream()), "dwconv_pack_weights")
    return packed


def dwconv_tc_res_gelu(x, y, n_img, h, w, packed, bias, k):
    """Tensor-core depthwise: y = gelu(x + dwconv_k(x) + b), fp16 rows views [n_img*h*w, C], packed from dwconv_pack_weights."""
    _cuda(x, y, packed, bias)
    _rows(x, "x"); _rows(y, "y")
    Cc = x.shape[1]
    if x.shape != y.shape or x.shape[0] != n_img * h * w or x.dtype != torch.float16 or y.dtype != torch.float16:
        raise ValueError("dwconv_tc_res_gelu: fp16 rows views of equal shape expected")
    if packed.numel() != _lib.load().vfml_dwconv_packed_bytes(Cc, k) or bias.numel() != Cc:
        raise ValueError("dwconv_tc_res_gelu: packed weights / bias do not match C and k")
    _lib.check(_lib.load().vfml_dwconv_tc_res_gelu(_p(x), _p(y), n_img, h, w, Cc, x.stride(0), y.stride(0), _p(packed),
                                                   _p(_f32c(bias, "b")), k, _stream()), "dwconv_tc_res_gelu")
    return y


def gemm_splitk(a, w, out, n_splits, partial, res=None, alpha=1.0):
    """out = res + alpha * a @ w^T with the K range cut into n_splits equal parts that run as the batch dimension of ONE
    launch (more CTAs than row tiles: a [7040, 7040] x [7040, 128] aggregate has only 55 row tiles for 148 SMs),
    partial products in `partial` (fp32, >= n_splits * M * N elements), combined by splitk_reduce."""
    _cuda(a, w, out, partial, res)
    _rows(a, "a"); _rows(w, "w"); _rows(out, "out")
    M, K = a.shape
    N = w.shape[0]
    if K % n_splits or w.shape[1] != K or out.shape != (M, N) or partial.dtype != torch.float32 or partial.numel() < n_splits * M * N:
        raise ValueError("gemm_splitk: shape mismatch")
    Ks = K // n_splits
    a3 = torch.as_strided(a, (n_splits, M, Ks), (Ks, a.stride(0), 1), a.storage_offset())
    w3 = torch.as_strided(w, (n_splits, N, Ks), (Ks, w.stride(0), 1), w.storage_offset())
    p3 = partial.reshape(-1)[: n_splits * M * N].view(n_splits, M, N)
    gemm_batched(a3, w3, p3, alpha=alpha)
    if res is not None:
        _rows(res, "res")
    _lib.check(_lib.load().vfml_splitk_reduce(_p(p3), n_splits, M, N, _p(res), res.stride(0) if res is not None else 0, _p(out),
                                              out.stride(0), _dt(out), _stream()), "splitk_reduce")
    return out


def layer_norm_rows(x, weight, bias, eps):
    """LayerNorm over the last dimension of a contiguous [..., C] tensor (fp32 or fp16, fp32 statistics)."""
    _cuda(x, weight, bias)
    if not x.is_contiguous() or x.dtype != weight.dtype or x.dtype != bias.dtype:
        raise ValueError("layer_norm_rows: x must be contiguous, weight / bias of the same dtype")
    Cc = x.shape[-1]
    y = torch.empty_like(x)
    rows = x.numel() // Cc
    _lib.check(_lib.load().vfml_layer_norm_rows(_p(x), _p(y), rows, Cc, Cc, Cc, _p(weight), _p(bias), float(eps), _dt(x), _stream()),
               "layer_norm_rows")
    return y


def attention_hd32(q, k, v, n_heads):
    """Fused attention with head dimension 32 on fp16 CUDA tensors.  q [B, Nq, n_heads*32], k / v [B, Nk, n_heads*32] may be
    strided views (slices of a fused qkv / kv projection: unit stride along the last axis) -> o [B, Nq, n_heads*32]."""
    _cuda(q, k, v)
    B, Nq, Cc = q.shape
    Nk = k.shape[1]
    if (Cc != n_heads * 32 or k.shape != (B, Nk, Cc) or v.shape != k.shape or q.dtype != torch.float16 or k.dtype != q.dtype
            or v.dtype != q.dtype or q.stride(2) != 1 or k.stride(2) != 1 or v.stride(2) != 1):
        raise ValueError("attention_hd32: expected fp16 [B, N, heads*32] tensors with unit stride along the channels")
    o = torch.empty((B, Nq, Cc), dtype=torch.float16, device=q.device)
    _lib.check(_lib.load().vfml_attention_hd32(_p(q), _p(k), _p(v), _p(o), B, n_heads, Nq, Nk, q.stride(0), q.stride(1), k.stride(0),
                                               k.stride(1), v.stride(0), v.stride(1), o.stride(0), o.stride(1), 32 ** -0.5, _stream()),
               "attention_hd32")
    return o


def peg3x3(x, n_img, h, w, weight_tapmajor, bias):
    """Twins PEG: x + dwconv3x3(x) + b on a contiguous token-major tensor [n_img, h*w, C]; weight [9, C], same dtype."""
    _cuda(x, weight_tapmajor, bias)
    Cc = x.shape[-1]
    if not x.is_contiguous() or x.numel() != n_img * h * w * Cc or weight_tapmajor.shape != (9, Cc) or bias.numel() != Cc \
            or not (x.dtype == weight_tapmajor.dtype == bias.dtype):
        raise ValueError("peg3x3: shape / dtype mismatch")
    y = torch.empty_like(x)
    _lib.check(_lib.load().vfml_peg3x3(_p(x), _p(y), n_img, h, w, Cc, Cc, Cc, _p(weight_tapmajor.contiguous()), _p(bias.contiguous()),
                                       _dt(x), _stream()), "peg3x3")
    return y


def softmax_rows(src, dst):
    _cuda(src, dst)
    _rows(src, "src"); _rows(dst, "dst")
    if src.dtype != torch.float32 or src.shape != dst.shape:
        raise ValueError("softmax_rows: src must be fp32 and match dst")
    _lib.check(_lib.load().vfml_softmax_rows(_p(src), _p(dst), src.shape[0], src.shape[1], src.stride(0), dst.stride(0),
                                             _dt(dst), _stream()), "softmax_rows")
    return dst


def motion_glue(fcoords, bcoords, x400, f128, mf, w1, b1, n_frames, h, w, rot=0):
    """rot: ring position of the frames (slot f holds the window's frame (f - rot) mod n_frames), include/vfml.h."""
    _cuda(fcoords, bcoords, x400, f128, mf, w1, b1)
    _rows(x400, "x400"); _rows(f128, "f128"); _rows(mf, "mf")
    rows = n_frames * h * w
    if x400.shape != (rows, 400) or f128.shape != (rows, 128) or mf.shape != (rows, 128):
        raise ValueError("motion_glue: shape mismatch")
    if not (x400.dtype == f128.dtype == mf.dtype):
        raise ValueError("motion_glue: dtype mismatch")
    _lib.check(_lib.load().vfml_motion_glue(_p(_f32c(fcoords, "fc")), _p(_f32c(bcoords, "bc")), _p(x400), x400.stride(0),
                                            _p(f128), f128.stride(0), _p(mf), mf.stride(0), 0, _p(_f32c(w1, "w1")),
                                            _p(_f32c(b1, "b1")), n_frames, h, w, int(rot), _dt(x400), _stream()), "motion_glue")


def coords_add(fcoords, bcoords, delta):
    _cuda(fcoords, bcoords, delta)
    _rows(delta, "delta")
    n, _, h, w = fcoords.shape
    if delta.dtype != torch.float32 or delta.shape != (n * h * w, 4):
        raise ValueError("coords_add: delta must be fp32 [n*h*w, 4]")
    _lib.check(_lib.load().vfml_coords_add(_p(_f32c(fcoords, "fc")), _p(_f32c(bcoords, "bc")), _p(delta), delta.stride(0),
                                           n, h * w, _stream()), "coords_add")


def coords_grid(n, h, w, device):
    out = torch.empty((n, 2, h, w), dtype=torch.float32, device=device)
    _cuda(out)
    _lib.check(_lib.load().vfml_coords_grid(_p(out), n, h, w, _stream()), "coords_grid")
    return out


def rows_fill(view, values):
    _cuda(view, values)
    _rows(view, "view")
    _lib.check(_lib.load().vfml_rows_fill(_p(view), view.stride(0), 0, view.shape[1], view.shape[0],
                                          _p(_f32c(values, "values")), _dt(view), _stream()), "rows_fill")


def im2col3x3(x, out, n_img, h, w):
    _cuda(x, out)
    _rows(x, "x"); _rows(out, "out")
    Cc = x.shape[1]
    if out.shape != (x.shape[0], 9 * Cc) or x.dtype != out.dtype:
        raise ValueError("im2col3x3: shape mismatch")
    _lib.check(_lib.load().vfml_im2col3x3(_p(x), _p(out), n_img, h, w, Cc, x.stride(0), out.stride(0), _dt(x), _stream()),
               "im2col3x3")
    return out


# ------------------------------------------------------------------ device guard
def _first_cuda_device(args):
    for a in args:
        if isinstance(a, torch.Tensor):
            if a.is_cuda:
                return a.device
        elif isinstance(a, TiledPyramid):
            return a.levels[0].device
        elif isinstance(a, torch.device) and a.type == "cuda":
            return a
        elif isinstance(a, str) and a.startswith("cuda"):
            return torch.device(a)
        elif isinstance(a, (list, tuple)) and a and isinstance(a[0], torch.Tensor) and a[0].is_cuda:
            return a[0].device
    return None


def _on_tensor_device(fn):
    """Kernels are launched on the current stream of the CURRENT device; run `fn` with the tensors' device current, so
    that a processor built for 'cuda:1' works without the caller calling torch.cuda.set_device first."""
    @functools.wraps(fn)
    def wrapper(*args, **kw):
        dev = _first_cuda_device(args) or _first_cuda_device(tuple(kw.values()))
        if dev is None or dev.index is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kw)
        with torch.cuda.device(dev):
            return fn(*args, **kw)
    return wrapper


for _name, _fn in list(globals().items()):
    if callable(_fn) and getattr(_fn, "__module__", None) == __name__ and not _name.startswith("_") \
            and _name not in ("pyramid_shapes", "corr_lookup_perm", "TiledPyramid"):
        globals()[_name] = _on_tensor_device(_fn)
del _name, _fn
