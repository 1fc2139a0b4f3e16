"""CPU restatement of the flow post-pass: gamedev / motion-vector RG8 encoding and TAA
(oracle; test infrastructure only).  numpy, same dtypes as the reference.

Pinned by ``tests/golden/post_*.npz`` (outputs of the unmodified reference classes
``encoding/flow_encoders.py`` and ``effects/taa_processor.py``).
"""
import numpy as np


def _to_u8(rgb):
    """flow_encoders.py:115-117: scale, scrub NaN/inf, truncate to uint8."""
    x = rgb * 255
    x = np.nan_to_num(x, nan=0.0, posinf=255.0, neginf=0.0)
    return x.astype(np.uint8)


def _rg_image(enc):
    h, w = enc.shape[:2]
    rgb = np.zeros((h, w, 3), dtype=np.float32)
    rgb[:, :, 0] = enc[:, :, 0]
    rgb[:, :, 1] = enc[:, :, 1]
    return rgb


def encode_gamedev(flow, width, height, scale=200.0, clamp=20.0):
    """GamedevFlowEncoder.encode. flow_encoders.py:90-117.  All arithmetic stays fp32
    (python-float scalars do not promote a float32 array)."""
    f = flow.astype(np.float32, copy=True)
    f[:, :, 0] /= width
    f[:, :, 1] /= height
    f *= scale
    c = np.clip(f, -clamp, clamp)
    e = np.clip((c + clamp) / (2 * clamp), 0, 1)
    return _to_u8(_rg_image(e))


def encode_mv_rg8(flow, clamp=64.0):
    """MotionVectorsRG8FlowEncoder.encode. flow_encoders.py:131-153"""
    c = np.clip(flow, -clamp, clamp)
    e = np.clip((c + clamp) / (2 * clamp), 0, 1)
    return _to_u8(_rg_image(e))


def decode_mv_rg8(img, clamp=64.0):
    """MotionVectorsRG8FlowEncoder.decode. flow_encoders.py:155-177"""
    n = img.astype(np.float32) / 255.0
    return (n[:, :, :2] * 2 * clamp) - clamp


class TAAOracle:
    """TAAProcessor with per-sequence history. effects/taa_processor.py:28-147.

    dtype behaviour follows the reference: first frame float32; with flow, the int64
    ``np.mgrid`` promotes coordinates (and from then on the history) to float64
    (SURVEY Q10)."""

    def __init__(self, alpha=0.1, sigma_color=25.0):
        self.alpha = alpha
        self.sigma = sigma_color
        self.history = {}

    def apply(self, frame, flow=None, alpha=None, use_flow=True, use_bilateral=True, seq="default"):
        a = self.alpha if alpha is None else alpha
        prev = self.history.get(seq)
        if prev is None:                                   # taa_processor.py:70-74
            out = frame.astype(np.float32)
        else:
            cur = frame.astype(np.float32)
            if not use_flow or flow is None:               # :78-80
                out = a * cur + (1 - a) * prev
            else:
                out = self._flow_taa(cur, flow, prev, a, use_bilateral)
        self.history[seq] = out
        return out

    def _flow_taa(self, cur, flow, prev, a, bilateral):
        # taa_processor.py:111-139
        h, w = cur.shape[:2]
        ys, xs = np.mgrid[0:h, 0:w]
        px = xs + flow[:, :, 0]
        py = ys + flow[:, :, 1]
        px = np.nan_to_num(px, nan=0.0, posinf=w - 1, neginf=0.0)
        py = np.nan_to_num(py, nan=0.0, posinf=h - 1, neginf=0.0)
        px = np.clip(px, 0, w - 1)
        py = np.clip(py, 0, h - 1)
        rep = self._bilateral(prev, px, py, cur) if bilateral else self._bilinear(prev, px, py)
        return a * cur + (1 - a) * rep

    def _bilateral(self, img, px, py, cur):
        # taa_processor.py:167-216
        h, w = img.shape[:2]
        x0 = np.clip(np.floor(px).astype(int), 0, w - 2)
        y0 = np.clip(np.floor(py).astype(int), 0, h - 2)
        x1, y1 = x0 + 1, y0 + 1
        wx = (px - x0)[..., None]
        wy = (py - y0)[..., None]
        nb = (img[y0, x0], img[y0, x1], img[y1, x0], img[y1, x1])
        lum = np.mean(cur, axis=2)
        s2 = self.sigma ** 2 * 0.1

        def cw(p):
            d = lum - np.mean(p, axis=2)
            return np.exp(-d ** 2 / (2 * s2 + 1e-6))[..., None]

        ws = ((1 - wx) * (1 - wy) * cw(nb[0]), wx * (1 - wy) * cw(nb[1]),
              (1 - wx) * wy * cw(nb[2]), wx * wy * cw(nb[3]))
        tot = ws[0] + ws[1] + ws[2] + ws[3]
        tot = np.where(tot == 0, 1e-6, tot)
        return (nb[0] * ws[0] + nb[1] * ws[1] + nb[2] * ws[2] + nb[3] * ws[3]) / tot

    def _bilinear(self, img, px, py):
        # taa_processor.py:233-262 (result buffer is float32)
        h, w = img.shape[:2]
        x0 = np.floor(px).astype(int)
        x1 = np.minimum(x0 + 1, w - 1)
        y0 = np.floor(py).astype(int)
        y1 = np.minimum(y0 + 1, h - 1)
        x0, x1 = np.clip(x0, 0, w - 1), np.clip(x1, 0, w - 1)
        y0, y1 = np.clip(y0, 0, h - 1), np.clip(y1, 0, h - 1)
        wx, wy = px - x0, py - y0
        out = np.zeros_like(img, dtype=np.float32)
        for c in range(img.shape[2]):
            out[:, :, c] = (img[y0, x0, c] * (1 - wx) * (1 - wy) + img[y0, x1, c] * wx * (1 - wy)
                            + img[y1, x0, c] * (1 - wx) * wy + img[y1, x1, c] * wx * wy)
        return out


# ----------------------------------------------------------------------------
# remaining encoders / decoders          encoding/flow_encoders.py
# ----------------------------------------------------------------------------
def encode_mv_rgb8(flow, clamp=32.0):
    """MotionVectorsRGB8FlowEncoder.encode, module-level ``method = 'rgb+'`` (flow_encoders.py:10,242-262,292-295):
    d = flow / clamp, renormalised to unit length where |d| > 1; R, G = (clip(d, -1, 1) + 1) / 2,
    B = sqrt(1 - dx^2 - dy^2) (NaN where rounding made the radicand negative -> 0).  fp32 throughout."""
    with np.errstate(invalid="ignore", divide="ignore"):
        dx = (flow[:, :, 0] / clamp).astype(np.float32)
        dy = (flow[:, :, 1] / clamp).astype(np.float32)
        ln = np.sqrt(dx ** 2 + dy ** 2)
        m = ln > 1
        dx[m] = dx[m] / ln[m]
        dy[m] = dy[m] / ln[m]
        corr = np.sqrt(1 - dx ** 2 - dy ** 2)
        rgb = np.zeros(flow.shape[:2] + (3,), dtype=np.float32)
        rgb[:, :, 0] = (np.clip(dx, -1, 1) + 1) / 2
        rgb[:, :, 1] = (np.clip(dy, -1, 1) + 1) / 2
        rgb[:, :, 2] = corr
        return _to_u8(rgb)


def decode_mv_rgb8(img, clamp=32.0):
    """MotionVectorsRGB8FlowEncoder.decode, 'rgb+' (flow_encoders.py:308,340-347,362-369)."""
    n = img.astype(np.float32) / 255.0
    dx = n[:, :, 0] * 2 - 1
    dy = n[:, :, 1] * 2 - 1
    ln = np.sqrt(dx ** 2 + dy ** 2 + n[:, :, 2] ** 2)
    mag = 1 / ln * clamp
    out = np.zeros(img.shape[:2] + (2,), dtype=np.float32)
    out[:, :, 0] = dx * mag
    out[:, :, 1] = dy * mag
    return out


def _fnma1(a, b):
    """fl32(1 - a * b) with ONE rounding (a fused multiply-add): the float64 product of two float32 is exact."""
    return (1.0 - a.astype(np.float64) * b.astype(np.float64)).astype(np.float32)


def hsv2rgb_u8(hsv):
    """cv2.cvtColor(hsv_u8, cv2.COLOR_HSV2RGB) restated.  Third-party arithmetic (opencv-python-headless 4.13 here,
    imgproc color_hsv HSV2RGB_b/_f; source not in the reference tree): 8-bit HSV with H in [0, 180] goes through the
    float path h' = H * (6/180), s = S * (1/255), v = V * (1/255); sector = floor(h') mod 6, f = h' - floor(h');
    tab = [v, v(1-s), v * fma(-s, f, 1), v * fma(-s, 1-f, 1)]; channel = tab[.] * 255, TRUNCATED for the pixels a row's
    32-pixel SIMD blocks cover (x < 32 * (W // 32)) and ROUNDED half-to-even in the scalar tail of the row.  The fused
    multiply-adds and the two roundings are what the shipped x86-64 (AVX2/FMA dispatch) binary does: determined here
    by exhaustive comparison with cv2 over every (H, S) at V = 255 and a V sweep, for W = 256 (all SIMD) and W = 1 (all
    tail), 0 mismatches; re-checked by tests/test_oracle_golden.py whenever cv2 is importable."""
    f32 = np.float32
    h = hsv[..., 0].astype(f32) * f32(6.0 / 180.0)
    s = hsv[..., 1].astype(f32) * f32(1.0 / 255.0)
    v = hsv[..., 2].astype(f32) * f32(1.0 / 255.0)
    sec = np.floor(h).astype(np.int32)
    f = h - sec.astype(f32)
    sec = np.where(sec >= 6, sec - 6, sec)
    one = f32(1.0)
    tab = np.stack([v, v * (one - s), v * _fnma1(s, f), v * _fnma1(s, one - f)], -1)
    # sector -> (b, g, r) indices into tab
    idx = np.array([[1, 3, 0], [1, 0, 2], [3, 0, 1], [0, 2, 1], [0, 1, 3], [2, 1, 0]], np.int64)[sec]
    bgr = np.take_along_axis(tab, idx, -1)
    bgr = np.where((s == 0)[..., None], v[..., None], bgr)
    x = bgr * f32(255.0)
    W = hsv.shape[-2]
    simd = (np.arange(W) < 32 * (W // 32)).reshape((1,) * (hsv.ndim - 2) + (W, 1))
    out = np.where(simd, np.trunc(x), np.rint(x)).clip(0, 255).astype(np.uint8)
    return out[..., ::-1].copy()


def encode_hsv(flow):
    """HSVFlowEncoder.encode (flow_encoders.py:38-67): hue = (atan2(y, x) + pi) / 2pi * 180 truncated, saturation =
    |flow| / max|flow| * 255 truncated, value 255, then OpenCV HSV -> RGB."""
    f = np.nan_to_num(flow, nan=0.0, posinf=1.0, neginf=-1.0)
    mag = np.sqrt(f[:, :, 0] ** 2 + f[:, :, 1] ** 2)
    ang = np.arctan2(f[:, :, 1], f[:, :, 0])
    hue = np.clip((ang + np.pi) / (2 * np.pi) * 180, 0, 180).astype(np.uint8)
    mx = np.max(mag) if mag.size else 0
    sat = (mag / mx * 255).astype(np.uint8) if mx > 0 else np.zeros_like(mag, dtype=np.uint8)
    val = np.full_like(mag, 255, dtype=np.uint8)
    return hsv2rgb_u8(np.stack([hue, sat, val], axis=2))


def _colorwheel():
    """torchvision.utils._make_colorwheel (Baker et al. colour wheel, 55 entries)."""
    RY, YG, GC, CB, BM, MR = 15, 6, 4, 11, 13, 6
    cw = np.zeros((RY + YG + GC + CB + BM + MR, 3), np.float32)
    fl = lambda n: np.floor(np.float32(255) * np.arange(n, dtype=np.float32) / np.float32(n))
    c = 0
    cw[0:RY, 0] = 255; cw[0:RY, 1] = fl(RY); c += RY
    cw[c:c + YG, 0] = 255 - fl(YG); cw[c:c + YG, 1] = 255; c += YG
    cw[c:c + GC, 1] = 255; cw[c:c + GC, 2] = fl(GC); c += GC
    cw[c:c + CB, 1] = 255 - fl(CB); cw[c:c + CB, 2] = 255; c += CB
    cw[c:c + BM, 2] = 255; cw[c:c + BM, 0] = fl(BM); c += BM
    cw[c:c + MR, 2] = 255 - fl(MR); cw[c:c + MR, 0] = 255
    return cw


def encode_torchvision(flow):
    """TorchvisionFlowEncoder.encode (flow_encoders.py:399-427): torchvision.utils.flow_to_image (third party,
    torchvision 0.26 here; restated from its source: normalise by max norm + eps, colour-wheel interpolation,
    floor(255 col)), followed by the reference's own ``(img * 255).astype(np.uint8)`` on the ALREADY uint8 image
    (:425), which wraps modulo 256: out = (-img) mod 256.  fp32 throughout."""
    f32 = np.float32
    u, v = flow[:, :, 0].astype(f32), flow[:, :, 1].astype(f32)
    mx = np.sqrt(u ** 2 + v ** 2).max()
    den = f32(mx) + f32(np.finfo(np.float32).eps)
    u, v = u / den, v / den
    norm = np.sqrt(u ** 2 + v ** 2)
    a = np.arctan2(-v, -u) / f32(np.pi)
    fk = (a + f32(1)) / f32(2) * f32(54)
    k0 = np.floor(fk).astype(np.int64)
    k1 = k0 + 1
    k1[k1 == 55] = 0
    fr = fk - k0.astype(f32)
    cw = _colorwheel()
    img = np.zeros(flow.shape[:2] + (3,), np.uint8)
    for c in range(3):
        col0 = cw[k0, c] / f32(255.0)
        col1 = cw[k1, c] / f32(255.0)
        col = (f32(1) - fr) * col0 + fr * col1
        col = f32(1) - norm * (f32(1) - col)
        img[:, :, c] = np.floor(f32(255) * col).astype(np.uint8)
    return (img * 255).astype(np.uint8)


# ----------------------------------------------------------------------------
# cache formats, LOD pyramid, quality map
# ----------------------------------------------------------------------------
def flo_bytes(flow):
    """FlowFileHandler.save_flow_flo (storage/cache_manager.py:18-30): 'PIEH', width, height (little-endian u32), fp32 data."""
    import struct
    h, w = flow.shape[:2]
    return b"PIEH" + struct.pack("<I", w) + struct.pack("<I", h) + flow.astype(np.float32).tobytes()


def lod_pyramid(flow, num_lods=5):
    """LODGenerator.generate_lods (storage/cache_manager.py:77-161): each level is the mean of the 2x2 block's cells
    that exist (odd sizes are padded bottom / right with weight 0), times 0.5."""
    lods = [flow]
    cur = flow.astype(np.float32)
    for _ in range(1, num_lods):
        h, w = cur.shape[:2]
        ph, pw = h % 2, w % 2
        wgt = np.pad(np.ones((h, w), np.float32), ((0, ph), (0, pw)))
        cur = np.pad(cur, ((0, ph), (0, pw), (0, 0)))
        H2, W2 = cur.shape[0] // 2, cur.shape[1] // 2
        blk = cur.reshape(H2, 2, W2, 2, 2)
        wb = wgt.reshape(H2, 2, W2, 2)
        # np.sum over the (contiguous temporary) 2x2 block: sequential, row-major: ((a00 + a01) + a10) + a11
        tot = ((wb[:, 0, :, 0] + wb[:, 0, :, 1]) + wb[:, 1, :, 0]) + wb[:, 1, :, 1]
        out = np.zeros((H2, W2, 2), np.float32)
        for c in range(2):
            x = blk[..., c] * wb
            sm = ((x[:, 0, :, 0] + x[:, 0, :, 1]) + x[:, 1, :, 0]) + x[:, 1, :, 1]
            out[:, :, c] = np.where(tot > 0, sm / np.maximum(tot, 1) * np.float32(0.5), 0)
        lods.append(out)
        cur = out
    return lods


def resize_flow_bilinear(flow, h, w):
    """F.interpolate(mode='bilinear', align_corners=False) of a flow field followed by the vector rescale
    (correction_worker.py:182-185): src = max((dst + 0.5) * in/out - 0.5, 0), neighbours clamped at the last texel."""
    f32 = np.float32
    fh, fw = flow.shape[:2]

    def axis(n_out, n_in):
        src = np.maximum((np.arange(n_out, dtype=f32) + f32(0.5)) * f32(n_in / n_out) - f32(0.5), 0).astype(f32)
        i0 = np.minimum(src.astype(np.int64), n_in - 1)
        i1 = np.minimum(i0 + 1, n_in - 1)
        l1 = (src - i0.astype(f32)).astype(f32)
        return i0, i1, f32(1) - l1, l1
    y0, y1, wy0, wy1 = axis(h, fh)
    x0, x1, wx0, wx1 = axis(w, fw)
    f = flow.astype(f32)
    top = f[y0][:, x0] * wx0[None, :, None] + f[y0][:, x1] * wx1[None, :, None]
    bot = f[y1][:, x0] * wx0[None, :, None] + f[y1][:, x1] * wx1[None, :, None]
    out = top * wy0[:, None, None] + bot * wy1[:, None, None]
    out[..., 0] *= f32(w / fw)
    out[..., 1] *= f32(h / fh)
    return out.astype(f32)


def quality_frame(frame1, frame2, flow, threshold):
    """generate_quality_frame_gpu (correction_worker.py:175-208): warp-compare colour similarity (mean of
    RGB-distance, mean-abs and cosine similarities) -> green (good) / red (bad) / pure red (out of bounds); a flow of
    another size is first resized bilinearly (:182-185)."""
    f32 = np.float32
    h, w = frame1.shape[:2]
    if flow.shape[:2] != (h, w):
        flow = resize_flow_bilinear(flow, h, w)
    a = frame1.astype(f32) / f32(255.0)
    b = frame2.astype(f32) / f32(255.0)
    ys, xs = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    tx = (xs - flow[..., 0]).astype(f32)
    ty = (ys - flow[..., 1]).astype(f32)
    oob = (tx < 0) | (tx >= w) | (ty < 0) | (ty >= h)
    with np.errstate(invalid="ignore"):
        txi = np.clip(np.nan_to_num(tx, nan=0.0, posinf=0.0, neginf=0.0).astype(np.int64), 0, w - 1)
        tyi = np.clip(np.nan_to_num(ty, nan=0.0, posinf=0.0, neginf=0.0).astype(np.int64), 0, h - 1)
    s = b[tyi, txi]
    d = a - s
    rgb_sim = f32(1.0) - np.sqrt((d ** 2).sum(-1)) / f32(1.732)
    abs_sim = f32(1.0) - np.abs(d).mean(-1)
    na = np.maximum(np.sqrt((a ** 2).sum(-1)), f32(1e-8))
    ns = np.maximum(np.sqrt((s ** 2).sum(-1)), f32(1e-8))
    cos = ((a / na[..., None]) * (s / ns[..., None])).sum(-1)
    sim = (rgb_sim + abs_sim + (cos + f32(1.0)) / f32(2.0)) / f32(3.0)
    green = np.clip((sim - f32(0.5)) * f32(2.0), 0, 1)
    red = np.clip(f32(1.0) - sim, 0, 1)
    good = sim > threshold
    q = np.zeros((h, w, 3), f32)
    q[..., 1] = np.where(good, green, 0)
    q[..., 0] = np.where(good, 0, red)
    q[oob] = (1.0, 0.0, 0.0)
    return (q * f32(255)).astype(np.uint8)
