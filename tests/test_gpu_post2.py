"""Remaining encoders / decoders, LOD pyramid and quality map through the C ABI against the outputs of
the unmodified reference (tests/golden/post2.npz) and against the numpy oracle on large adversarial inputs.  pytest -m gpu."""
import numpy as np
import pytest
import torch

import cases
from oracle import post as opost
from test_oracle_golden import _post2_inputs

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_rgb8_hsv_torchvision_wrappers_bit_exact(golden):
    from video_flow_ml_b200.encoding import (FlowEncoderFactory, HSVFlowEncoder, MotionVectorsRGB8FlowEncoder,
                                             TorchvisionFlowEncoder, decode_motion_vectors, encode_flow, encode_motion_vectors)
    g = golden("post2.npz")
    ff, fin, _, _, _ = _post2_inputs()
    h, w = ff.shape[:2]
    assert set(FlowEncoderFactory.get_available_formats()) == {"hsv", "gamedev", "torchvision", "motion-vectors-rg8",
                                                               "motion-vectors-rgb8"}
    assert np.array_equal(MotionVectorsRGB8FlowEncoder().encode(ff.copy(), w, h), g["rgb8"])
    assert np.array_equal(FlowEncoderFactory.create_encoder("motion-vectors-rgb8", clamp_range=8.0).encode(ff, w, h), g["rgb8_c8"])
    assert np.array_equal(MotionVectorsRGB8FlowEncoder().decode(g["rgb8"]), g["rgb8_decoded"])
    assert np.array_equal(encode_motion_vectors(ff.copy()), g["mv_default"])                 # default variant: rgb8 (:502)
    assert np.array_equal(decode_motion_vectors(g["mv_default"]), g["mv_default_decoded"])
    assert np.array_equal(HSVFlowEncoder().encode(ff.copy(), w, h), g["hsv"])
    assert np.array_equal(encode_flow(fin.copy(), 61, 45, "hsv"), g["hsv_fin"])
    assert np.array_equal(HSVFlowEncoder().encode(np.zeros((5, 7, 2), np.float32), 7, 5), g["hsv_zero"])
    assert np.array_equal(TorchvisionFlowEncoder().encode(fin.copy(), 61, 45), g["tv"])
    dv = MotionVectorsRGB8FlowEncoder().encode(dev(ff), w, h)                                # device in -> device out
    assert dv.is_cuda and np.array_equal(dv.cpu().numpy(), g["rgb8"])


def test_encoders_full_size_vs_oracle():
    """1080p adversarial field (large values, NaN / inf, exact clamp boundaries, ties): bit-exact for rgb8 / decode.
    HSV and torchvision go through atan2; the kernels use the correctly rounded value, numpy (SIMD) / torch (Sleef) are
    within 1 ulp of it, so a pixel whose hue or colour-wheel position sits within that ulp of an integer boundary can land
    on the other side: at most 1e-5 of the pixels may differ (measured: 1 pixel of 2,073,600 for HSV, 0 for torchvision)."""
    from video_flow_ml_b200 import ops
    r = np.random.default_rng(5)
    f = (30 * r.standard_normal((1080, 1920, 2))).astype(np.float32)
    f[::97, ::89] = [np.nan, 1.0]
    f[::101, ::83] = [np.inf, -np.inf]
    f[5] = 0.0
    f[6, :, 1] = 0.0
    f[7, :, 0] = 0.0
    f[8] = 32.0
    d = dev(f)
    with np.errstate(all="ignore"):
        enc = ops.encode_mv_rgb8(d, 32.0).cpu().numpy()
        assert np.array_equal(enc, opost.encode_mv_rgb8(f, 32.0))
        assert np.array_equal(ops.decode_mv_rgb8(dev(enc), 32.0).cpu().numpy(), opost.decode_mv_rgb8(enc, 32.0))
        for name, got, want in (("hsv", ops.encode_hsv(d).cpu().numpy(), opost.encode_hsv(f)),
                                ("torchvision", ops.encode_torchvision(dev(np.nan_to_num(f, nan=0.0, posinf=9.0, neginf=-9.0))).cpu().numpy(),
                                 opost.encode_torchvision(np.nan_to_num(f, nan=0.0, posinf=9.0, neginf=-9.0)))):
            bad = (got != want).any(-1)
            print(f"{name}: {int(bad.sum())} of {bad.size} pixels differ")
            assert bad.mean() <= 1e-5, (name, int(bad.sum()))
    nowrap = ops.encode_torchvision(dev(np.nan_to_num(f, nan=0.0, posinf=9.0, neginf=-9.0)), ref_wrap=False).cpu().numpy()
    assert np.array_equal((nowrap.astype(np.int32) * 255 % 256).astype(np.uint8),
                          ops.encode_torchvision(dev(np.nan_to_num(f, nan=0.0, posinf=9.0, neginf=-9.0))).cpu().numpy())


def test_lod_pyramid_bit_exact(golden):
    from video_flow_ml_b200 import ops
    from video_flow_ml_b200.storage import LODGenerator
    g = golden("post2.npz")
    _, fin, fe, _, _ = _post2_inputs()
    for name, f in (("lod_odd", fin), ("lod_even", fe)):
        lods = LODGenerator.generate_lods(f, 5)
        assert lods[0] is f and len(lods) == 5
        for i in range(1, 5):
            assert lods[i].dtype == np.float32 and np.array_equal(lods[i], g[f"{name}_{i}"]), (name, i)
    # ragged full-size field, 7 levels (two chained launches), against the oracle
    r = np.random.default_rng(6)
    f = (20 * r.standard_normal((437, 1023, 2))).astype(np.float32)
    got = ops.lod_pyramid(dev(f), 7)
    want = opost.lod_pyramid(f, 7)
    assert [tuple(t.shape) for t in got] == [w.shape for w in want]
    for a, b in zip(got[1:], want[1:]):
        assert np.array_equal(a.cpu().numpy(), b)
    one = ops.lod_pyramid(dev(f[:1, :1]), 3)
    assert [tuple(t.shape) for t in one] == [(1, 1, 2)] * 3 and np.array_equal(one[2].cpu().numpy(), f[:1, :1] * 0.25)


def test_quality_frame(golden):
    from video_flow_ml_b200.quality import generate_quality_frame_gpu
    g = golden("post2.npz")
    _, _, _, frames, fl = _post2_inputs()
    assert np.array_equal(generate_quality_frame_gpu(frames[0], frames[1], fl, "cuda", 0.7), g["quality"])
    assert np.array_equal(generate_quality_frame_gpu(frames[0], frames[0], np.zeros_like(fl), "cuda", 0.7), g["quality_same"])
    small = np.ascontiguousarray(fl[::2, ::2][:16, :23])
    assert np.array_equal(generate_quality_frame_gpu(frames[0], frames[1], small, "cuda", 0.7), g["quality_resized"])
    r = np.random.default_rng(8)
    a, b = (r.integers(0, 256, (436, 1024, 3)).astype(np.uint8) for _ in range(2))
    b[::3] = a[::3]
    f = (6 * r.standard_normal((436, 1024, 2))).astype(np.float32)
    f[10, 10] = [np.nan, 0.0]
    got = generate_quality_frame_gpu(dev(a), dev(b), dev(f), "cuda", 0.6)
    assert got.is_cuda and np.array_equal(got.cpu().numpy(), opost.quality_frame(a, b, f, 0.6))


def test_taa_float_frames_and_wrappers(golden):
    """Float current frames keep their fractions (reference: current_frame.astype(np.float32), taa_processor.py:64);
    TAAComparisonProcessor / apply_taa_effect mirror :322-425."""
    from video_flow_ml_b200.effects import TAAComparisonProcessor, TAAProcessor, apply_taa_effect
    frames, flows = cases.taa_case()
    ff = [f.astype(np.float32) + 0.25 for f in frames]
    t, o = TAAProcessor(alpha=0.1), opost.TAAOracle(alpha=0.1)
    prev = None
    for i, f in enumerate(ff):
        got = t.apply_taa(f, prev, alpha=0.1, sequence_id="s")
        want = o.apply(f, prev, alpha=0.1, seq="s")
        assert got.dtype == want.dtype and np.abs(got.astype(np.float64) - want).max() <= 1e-3, i
        prev = flows[i].copy()
    g = golden("post.npz")
    c = TAAComparisonProcessor(alpha=0.1)
    prev = None
    for i, f in enumerate(frames):
        a, b = c.apply_comparison(f, prev, alpha=0.1)
        assert np.abs(a.astype(np.float64) - g[f"taa_bilateral_{i}"]).max() <= 1e-3
        assert np.abs(b.astype(np.float64) - g[f"taa_simple_{i}"]).max() <= 1e-3
        prev = flows[i].copy()
    assert np.array_equal(apply_taa_effect(frames[0]), frames[0].astype(np.float32))
