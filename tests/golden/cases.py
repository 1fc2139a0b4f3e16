"""Seeded input generators shared by ``make_golden.py`` (which feeds them to the unmodified
reference) and by the tests (which feed them to the oracle and to the CUDA path).

numpy PCG64 only, so the inputs are bit-identical on every machine.
"""
import numpy as np


def rng(seed):
    return np.random.default_rng(seed)


def synthetic_clip(n_frames, height, width, seed=1234, max_shift=24.0):
    """uint8 RGB frames: multi-octave smooth noise texture translated by a smooth,
    frame-dependent motion plus 2 % salt noise."""
    r = rng(seed)
    H, W = height + 128, width + 128
    tex = np.zeros((H, W, 3), np.float32)
    for octave in range(5):
        s = 2 ** (octave + 2)
        g = r.standard_normal((H // s + 2, W // s + 2, 3)).astype(np.float32)
        up = np.kron(g, np.ones((s, s, 1), np.float32))[:H, :W]
        tex += up / (octave + 1)
    # cheap separable blur so that octaves are smooth
    k = np.array([1, 4, 6, 4, 1], np.float32) / 16
    for ax in (0, 1):
        tex = sum(np.roll(tex, i - 2, axis=ax) * k[i] for i in range(5))
    tex = (tex - tex.min()) / (tex.max() - tex.min())
    frames = []
    for t in range(n_frames):
        dx = int(round(max_shift * np.sin(0.7 * t) * 0.5 + 3 * t))
        dy = int(round(max_shift * np.cos(0.4 * t) * 0.25 + 2 * t))
        dx = int(np.clip(64 + dx, 0, 128))
        dy = int(np.clip(64 + dy, 0, 128))
        f = tex[dy:dy + height, dx:dx + width].copy()
        salt = r.random((height, width, 1)) < 0.02
        f = np.where(salt, r.random((height, width, 3)).astype(np.float32), f)
        frames.append((f * 255).astype(np.uint8))
    return frames


def lookup_case(seed=7, B=2, C=256, h=16, w=24):
    r = rng(seed)
    f1 = r.standard_normal((B, C, h, w)).astype(np.float32)
    f2 = r.standard_normal((B, C, h, w)).astype(np.float32)
    ys, xs = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    grid = np.stack([xs, ys], 0).astype(np.float32)[None].repeat(B, 0)
    coords = grid + (6.0 * r.standard_normal((B, 2, h, w))).astype(np.float32)
    coords[0, :, 0, 0] = grid[0, :, 0, 0]               # exact integer coordinate
    coords[0, :, 1, 1] = [-30.0, 5.0]                    # far out of range
    coords[0, :, 2, 2] = [w - 1.0, h - 1.0]              # last texel exactly
    coords[0, :, 3, 3] = [w - 0.5, h - 0.5]              # straddles the border
    return f1, f2, coords


def update_case(seed=11, Fc=3, h=16, w=24):
    r = rng(seed)
    g = lambda *s: r.standard_normal(s).astype(np.float32)
    d = dict(net=np.tanh(g(Fc, 128, h, w)), inp=np.maximum(g(Fc, 128, h, w), 0),
             fcorr=2 * g(Fc, 324, h, w), bcorr=2 * g(Fc, 324, h, w),
             fflow=3 * g(Fc, 2, h, w), bflow=3 * g(Fc, 2, h, w), mhs=g(Fc, 48, h, w))
    return d


def upsample_case(seed=13, N=2, h=9, w=11):
    r = rng(seed)
    flow = (4 * r.standard_normal((N, 2, h, w))).astype(np.float32)
    mask = (3 * r.standard_normal((N, 576, h, w))).astype(np.float32)
    return flow, mask


def flow_field_case(seed=17, h=37, w=53):
    """fp32 flow [h,w,2] with large values, exact clamp boundaries, NaN and +-inf."""
    r = rng(seed)
    f = (12 * r.standard_normal((h, w, 2))).astype(np.float32)
    f[0, 0] = [np.nan, np.inf]
    f[0, 1] = [-np.inf, 0.0]
    f[0, 2] = [64.0, -64.0]
    f[0, 3] = [w * 0.1, -h * 0.1]                         # exactly +-20 after gamedev scaling
    f[0, 4] = [1e-30, -1e-30]
    f[1, :, 0] = np.linspace(-80, 80, w, dtype=np.float32)
    return f


def taa_case(seed=19, n=4, h=21, w=30):
    r = rng(seed)
    frames = [r.integers(0, 256, (h, w, 3)).astype(np.uint8) for _ in range(n)]
    flows = [(3 * r.standard_normal((h, w, 2))).astype(np.float32) for _ in range(n)]
    flows[1][0, 0] = [np.nan, -np.inf]
    flows[1][0, 1] = [np.inf, 100.0]
    flows[2][5, 5] = [-100.0, -100.0]
    return frames, flows


def twins_case(seed=23, B=2, H=64, W=96):
    r = rng(seed)
    return r.uniform(-1, 1, (B, 3, H, W)).astype(np.float32)


def mof_case(seed=29, n_frames=6, H=124, W=188, as_float=True):
    """Clip whose size is NOT a multiple of 8 (exercises InputPadder 'sintel' mode).
    ``as_float``: float32 frames holding 0..255 values -- prepare_frame_sequence passes float
    frames through unscaled (videoflow_processor.py:155-157), so the network sees properly
    normalised images and the recurrence is input-driven.  uint8 frames take the /255 branch
    (double normalisation, near-constant features)."""
    fr = synthetic_clip(n_frames, H, W, seed=seed, max_shift=10.0)
    return [f.astype(np.float32) for f in fr] if as_float else fr


_TEX = {}


def _texture(height, width, seed):
    """The multi-octave smooth noise texture of synthetic_clip (same construction), cached per size."""
    key = (height, width, seed)
    if key not in _TEX:
        r = rng(seed)
        H, W = height + 128, width + 128
        tex = np.zeros((H, W, 3), np.float32)
        for octave in range(5):
            s = 2 ** (octave + 2)
            g = r.standard_normal((H // s + 2, W // s + 2, 3)).astype(np.float32)
            tex += np.kron(g, np.ones((s, s, 1), np.float32))[:H, :W] / (octave + 1)
        k = np.array([1, 4, 6, 4, 1], np.float32) / 16
        for ax in (0, 1):
            tex = sum(np.roll(tex, i - 2, axis=ax) * k[i] for i in range(5))
        _TEX.clear()
        _TEX[key] = (tex - tex.min()) / (tex.max() - tex.min())
    return _TEX[key]


def sliding_clip(start, stop, height, width, seed=1234, max_shift=24.0):
    """Frames [start, stop) of an endless synthetic clip (bench.py): like synthetic_clip, but frame t depends only on
    (seed, t) -- a bounded Lissajous translation of the texture plus 2 % salt noise drawn from a per-frame generator --
    so every rank of a sharded job can generate just its own range, and no two frames are ever equal."""
    tex = _texture(height, width, seed)
    frames = []
    for t in range(start, stop):
        dx = int(np.clip(64 + round(max_shift * 2.0 * np.sin(0.31 * t) + 12 * np.sin(0.047 * t)), 0, 128))
        dy = int(np.clip(64 + round(max_shift * 1.0 * np.cos(0.23 * t) + 8 * np.sin(0.061 * t)), 0, 128))
        f = tex[dy:dy + height, dx:dx + width]
        r = np.random.default_rng([seed, t])
        salt = r.random((height, width, 1), dtype=np.float32) < 0.02
        f = np.where(salt, r.random((height, width, 3), dtype=np.float32), f)
        frames.append((f * 255).astype(np.uint8))
    return frames
