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
