"""The polynomial GELU of the tensor-core epilogue (csrc/gelu_poly.cuh gelu_poly2), re-evaluated in numpy fp32 with the
coefficients parsed from the CUDA source, against the exact erf GELU: max abs error <= 3e-6 over [-12, 12]."""
import math
import os
import re

import numpy as np

SRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "video_flow_ml_b200", "csrc", "gelu_poly.cuh")


def test_gelu_polynomial_error():
    body = open(SRC).read().split("gelu_poly2(float2 x) {")[1].split("return")[0]
    coef = [np.float32(c) for c in re.findall(r"splat\((-?\d\.\d+e[+-]\d+)f\)", body)]
    assert len(coef) == 13
    x = np.linspace(-12, 12, 600001).astype(np.float32)
    xc = np.clip(x, -5, 5)
    s = (xc * xc * np.float32(0.08) - np.float32(1)).astype(np.float32)
    r = np.full_like(s, coef[0])
    for c in coef[1:]:
        r = (r * s + c).astype(np.float32)
    got = (x * (xc * r + np.float32(0.5)).astype(np.float32)).astype(np.float32)
    erf = np.vectorize(math.erf)
    want = 0.5 * x.astype(np.float64) * (1 + erf(x.astype(np.float64) / math.sqrt(2)))
    assert np.abs(got - want).max() <= 4e-6, np.abs(got - want).max()


def test_gelu_tanh_form_fit_error():
    """gelu_tanh2 (the MUFU variant): x/2 * (1 + tanh(xc * Q(xc^2))), xc = clip(x, +-7), with the coefficients parsed
    from the CUDA source and an exact tanh: fit error <= 4e-5 abs over [-12, 12].  (tanh.approx.f32 adds up to
    2^-11 relative on the device; the end-to-end effect is what tests/test_gpu_e2e.py bounds.)"""
    body = open(SRC).read().split("gelu_tanh2(float2 x) {")[1].split("return")[0]
    c2, c1, c0 = [np.float32(c) for c in re.findall(r"splat\((-?\d\.\d+e[+-]\d+)f\)", body)[:3]]
    x = np.linspace(-12, 12, 600001).astype(np.float32)
    xc = np.clip(x, -7, 7)
    x2 = (xc * xc).astype(np.float32)
    q = ((c2 * x2 + c1).astype(np.float32) * x2 + c0).astype(np.float32)
    th = np.tanh((xc * q).astype(np.float32)).astype(np.float32)
    got = (np.float32(0.5) * x * th + np.float32(0.5) * x).astype(np.float32)
    erf = np.vectorize(math.erf)
    want = 0.5 * x.astype(np.float64) * (1 + erf(x.astype(np.float64) / math.sqrt(2)))
    assert np.abs(got - want).max() <= 4e-5, np.abs(got - want).max()
