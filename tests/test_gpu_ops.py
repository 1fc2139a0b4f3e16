"""Parity of every libvfml kernel (called through the C ABI) against the oracle / golden fixtures.
Runs on the B200 box: pytest -m gpu."""
import numpy as np
import pytest
import torch

import cases
import fake_ops
from oracle import mof as omof
from oracle import post as opost

pytestmark = pytest.mark.gpu
T = torch.from_numpy


@pytest.fixture(scope="module")
def ops():
    from video_flow_ml_b200 import ops as o
    return o


def cu(a):
    return (T(a) if isinstance(a, np.ndarray) else a).cuda()


def close(got, want, atol, rtol=0.0):
    got, want = got.detach().double().cpu(), want.detach().double().cpu()
    err = (got - want).abs().max().item()
    assert err <= atol + rtol * want.abs().max().item(), f"max err {err:.3e} (ref max {want.abs().max().item():.3e})"


def rows_buf(n, c, dtype=torch.float32, pad=8):
    return torch.zeros((n, c + pad), dtype=dtype, device="cuda")[:, :c]


# ---------------------------------------------------------------- correlation volume + lookup
def test_volume_pyramid_and_lookup_golden(ops, golden):
    g = golden("corr_lookup.npz")
    f1, f2, coords = cases.lookup_case()
    pyr = ops.corr_volume_pyramid(cu(f1), cu(f2), 4)
    assert [tuple(p.shape) for p in pyr] == [(768, 1, 16, 24), (768, 1, 8, 12), (768, 1, 4, 6), (768, 1, 2, 3)]
    close(pyr[0].view(2, 16, 24, 1, 16, 24)[0, :2], T(g["vol0"]), 2e-5)
    close(pyr[3][:64], T(g["pyr3"]), 2e-5)
    out = ops.corr_lookup(pyr, cu(coords), 4)
    assert out.shape == (2, 324, 16, 24) and out.is_contiguous()
    close(out, T(g["lookup"]), 5e-5)
    rows = rows_buf(2 * 16 * 24, 324)
    ops.corr_lookup(pyr, cu(coords), 4, out=rows)
    close(rows.view(2, 16, 24, 324).permute(0, 3, 1, 2), T(g["lookup"]), 5e-5)


@pytest.mark.parametrize("n,h,w", [(1, 16, 17), (3, 20, 31), (2, 16, 40)])
def test_lookup_vs_oracle_ragged(ops, n, h, w):
    r = np.random.default_rng(5)
    f1 = r.standard_normal((n, 64, h, w)).astype(np.float32)
    f2 = r.standard_normal((n, 64, h, w)).astype(np.float32)
    pyr_o = omof.corr_pyramid(T(f1), T(f2))
    coords = omof.pixel_grid(n, h, w) + T((5 * r.standard_normal((n, 2, h, w))).astype(np.float32))
    want = omof.corr_lookup(pyr_o, coords)
    pyr = ops.corr_volume_pyramid(cu(f1), cu(f2), 4)
    for a, b in zip(pyr, pyr_o):
        close(a, b, 2e-5)
    close(ops.corr_lookup(pyr, coords.cuda(), 4), want, 5e-5)
    # radius 3 / 2 levels goes through the generic kernel
    want2 = omof.corr_lookup(pyr_o[:2], coords, radius=3)
    close(ops.corr_lookup(pyr[:2], coords.cuda(), 3), want2, 5e-5)
    # fp16 volume storage (tensor-core tier): compare against a lookup in the rounded volume
    pyr16 = [p.half() for p in pyr]
    want16 = omof.corr_lookup([p.half().float() for p in pyr_o], coords)
    got16 = ops.corr_lookup(pyr16, coords.cuda(), 4, out_dtype=torch.float16)
    close(got16, want16, 2e-2, 2e-3)


def test_lookup_empty_and_errors(ops):
    f = torch.zeros((0, 16, 8, 8), device="cuda")
    pyr = ops.corr_volume_pyramid(f, f, 4)
    out = ops.corr_lookup(pyr, torch.zeros((0, 2, 8, 8), device="cuda"), 4)
    assert out.shape == (0, 324, 8, 8)
    with pytest.raises(ValueError):
        ops.corr_lookup(pyr, torch.zeros((1, 2, 8, 8)), 4)          # CPU tensor: no fallback
    with pytest.raises(ValueError):
        ops.corr_volume_pyramid(torch.zeros((1, 16, 4, 4), device="cuda"), torch.zeros((1, 16, 4, 4), device="cuda"), 4)


# ---------------------------------------------------------------- upsample
def test_convex_upsample_golden(ops, golden):
    fl, mk = cases.upsample_case()
    want = T(golden("upsample.npz")["up"])
    close(ops.convex_upsample8(cu(fl), cu(mk)), want, 2e-5)
    n, _, h, w = fl.shape
    rows = rows_buf(n * h * w, 576)
    rows.copy_(cu(mk).permute(0, 2, 3, 1).reshape(n * h * w, 576))
    close(ops.convex_upsample8(cu(fl), rows), want, 2e-5)
    # cropped write into a larger frame (unpad + tile paste) and coords input
    frame = torch.full((n, 2, 100, 120), -7.0, device="cuda")
    grid = omof.pixel_grid(n, h, w)
    ops.convex_upsample8(cu(fl) + grid.cuda(), rows, out=frame[:, :, 10:10 + 60, 20:20 + 70], crop=(3, 5, 60, 70),
                         flow_is_coords=True)
    close(frame[:, :, 10:70, 20:90], want[:, :, 3:63, 5:75], 5e-5)
    assert (frame[:, :, :10] == -7).all() and (frame[:, :, :, 90:] == -7).all()


# ---------------------------------------------------------------- post-pass
def test_encoders_bit_exact(ops, golden):
    g = golden("post.npz")
    ff = cases.flow_field_case()
    h, w = ff.shape[:2]
    assert np.array_equal(ops.encode_gamedev(cu(ff), w, h).cpu().numpy(), g["gamedev"])
    assert np.array_equal(ops.encode_gamedev(cu(ff), 3840, 2160).cpu().numpy(), g["gamedev_4k"])
    assert np.array_equal(ops.encode_mv_rg8(cu(ff)).cpu().numpy(), g["rg8"])
    assert np.array_equal(ops.encode_mv_rg8(cu(ff), 16.0).cpu().numpy(), g["rg8_c16"])
    assert np.array_equal(ops.decode_mv_rg8(cu(g["rg8"])).cpu().numpy(), g["rg8_decoded"])


def test_encoders_bit_exact_full_size(ops):
    """4K frame of adversarial values against the numpy oracle: bit-exact."""
    r = np.random.default_rng(3)
    ff = (30 * r.standard_normal((2160, 3840, 2))).astype(np.float32)
    ff[::7, ::5] *= 1e-3
    got = ops.encode_gamedev(cu(ff), 3840, 2160).cpu().numpy()
    assert np.array_equal(got, opost.encode_gamedev(ff, 3840, 2160))
    assert np.array_equal(ops.encode_mv_rg8(cu(ff)).cpu().numpy(), opost.encode_mv_rg8(ff))


@pytest.mark.parametrize("name,mode", [("taa_bilateral", 2), ("taa_bilinear", 1), ("taa_simple", 0)])
def test_taa_golden(ops, golden, name, mode):
    g = golden("post.npz")
    frames, flows = cases.taa_case()
    hist = None
    prev = None
    for i, f in enumerate(frames):
        if hist is None:
            hist = ops.taa_init(cu(f))
        else:
            hist_f32 = (mode != 2) or i == 1
            hist = ops.taa(cu(f), cu(prev) if mode else None, hist, 0.1, 25.0, mode, hist_f32)
        want = g[f"{name}_{i}"]
        err = np.abs(hist.cpu().numpy() - want.astype(np.float64)).max()
        assert err <= 1e-3, (name, i, err)
        prev = flows[i].copy()


# ---------------------------------------------------------------- update-step building blocks vs the CPU double
@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-4), (torch.float16, 3e-2)])
@pytest.mark.parametrize("M,N,K", [(300, 128, 324), (257, 4, 192), (128, 486, 324), (64, 64, 7), (1000, 172, 600)])
def test_gemm_epilogues(ops, dt, tol, M, N, K):
    r = torch.Generator().manual_seed(M + N + K)
    a = torch.randn((M, K), generator=r).to(dt)
    w = (torch.randn((N, K), generator=r) / K ** 0.5).to(dt)
    bias, dww, dwb = torch.randn(N, generator=r), torch.randn(N, generator=r) * 0.3, torch.randn(N, generator=r) * 0.3
    res = torch.randn((M, N), generator=r).to(dt)
    for kw in (dict(), dict(bias=bias, act="gelu"), dict(bias=bias, res=res, act="gelu", dw=(dww, dwb)),
               dict(bias=bias, act="relu", alpha=100.0), dict(res=res, alpha=0.5)):
        want = torch.zeros((M, N), dtype=dt)
        fake_ops.gemm(a, w, want, **kw)
        got = rows_buf(M, N, dt)
        kwc = {k: (tuple(cu(x) for x in v) if isinstance(v, tuple) else cu(v) if torch.is_tensor(v) else v) for k, v in kw.items()}
        A = rows_buf(M, K, dt, pad=16); A.copy_(a)
        ops.gemm(A, cu(w), got, **kwc)
        close(got, want, tol, tol)
    out32 = rows_buf(M, N, torch.float32)
    ops.gemm(cu(a), cu(w), out32)
    close(out32, a.float() @ w.float().t(), tol, tol)


@pytest.mark.parametrize("dt,tol", [(torch.float32, 1e-4), (torch.float16, 2e-2)])
@pytest.mark.parametrize("k", [7, 15])
def test_dwconv(ops, dt, tol, k):
    r = torch.Generator().manual_seed(k)
    n, h, w, C = 2, 13, 21, 44
    x = torch.randn((n * h * w, C), generator=r).to(dt)
    wt = torch.randn((k * k, C), generator=r) / k
    b = torch.randn(C, generator=r)
    want = torch.zeros_like(x)
    fake_ops.dwconv_res_gelu(x, want, n, h, w, wt, b, k)
    X = rows_buf(n * h * w, C, dt); X.copy_(x)
    got = rows_buf(n * h * w, C, dt)
    ops.dwconv_res_gelu(X, got, n, h, w, cu(wt), cu(b), k)
    close(got, want, tol, tol)


def test_misc_update_ops(ops):
    r = torch.Generator().manual_seed(1)
    Fc, h, w = 3, 11, 17
    Np = Fc * h * w
    # layout round trip
    src = torch.randn((Fc, 37, h, w), generator=r)
    rows = rows_buf(Np, 37)
    ops.nchw_to_rows(cu(src), rows)
    close(rows, src.permute(0, 2, 3, 1).reshape(Np, 37), 0)
    back = torch.empty((Fc, 37, h, w), device="cuda")
    ops.rows_to_nchw(rows, back)
    close(back, src, 0)
    # softmax
    s = torch.randn((50, 1000), generator=r) * 5
    out = torch.empty((50, 1000), device="cuda")
    ops.softmax_rows(cu(s), out)
    close(out, torch.softmax(s, -1), 1e-6)
    # coords grid / add
    g = ops.coords_grid(Fc, h, w, "cuda")
    close(g, omof.pixel_grid(Fc, h, w), 0)
    fc, bc = g.clone(), g.clone()
    delta = torch.randn((Np, 4), generator=r)
    ops.coords_add(fc, bc, cu(delta))
    fcc, bcc = omof.pixel_grid(Fc, h, w), omof.pixel_grid(Fc, h, w)
    fake_ops.coords_add(fcc, bcc, delta)
    close(fc, fcc, 1e-6); close(bc, bcc, 1e-6)
    # motion glue
    x400 = torch.randn((Np, 400), generator=r)
    f128, mf = torch.zeros((Np, 128)), torch.zeros((Np, 128))
    w1, b1 = torch.randn((128, 4), generator=r), torch.randn(128, generator=r)
    fcd = fcc + 3 * torch.randn((Fc, 2, h, w), generator=r)
    bcd = bcc + 3 * torch.randn((Fc, 2, h, w), generator=r)
    X, F1, MF = rows_buf(Np, 400), rows_buf(Np, 128), rows_buf(Np, 128)
    X.copy_(x400)
    x0_400 = x400.clone()
    fake_ops.motion_glue(fcd, bcd, x400, f128, mf, w1, b1, Fc, h, w)
    ops.motion_glue(cu(fcd), cu(bcd), X, F1, MF, cu(w1), cu(b1), Fc, h, w)
    close(X, x400, 2e-5); close(F1, f128, 2e-5); close(MF[:, 124:], mf[:, 124:], 1e-6)
    # ring order (the engine keeps a sliding window's frames in place and rotates): frames rolled by `rot` slots give the
    # rolled result, on the test double (vs rot = 0) and on the kernel (vs the test double)
    for rot in range(1, Fc):
        roll_r = lambda t: torch.roll(t.reshape(Fc, h * w, -1), rot, 0).reshape(Np, -1)
        xr, fr, mr = roll_r(x0_400), torch.zeros((Np, 128)), torch.zeros((Np, 128))
        X.copy_(xr)
        fake_ops.motion_glue(torch.roll(fcd, rot, 0), torch.roll(bcd, rot, 0), xr, fr, mr, w1, b1, Fc, h, w, rot=rot)
        close(xr, roll_r(x400), 1e-6); close(fr, roll_r(f128), 1e-6)
        ops.motion_glue(cu(torch.roll(fcd, rot, 0)), cu(torch.roll(bcd, rot, 0)), X, F1, MF, cu(w1), cu(b1), Fc, h, w, rot=rot)
        close(X, xr, 2e-5); close(F1, fr, 2e-5); close(MF[:, 124:], mr[:, 124:], 1e-6)
    # rows fill, im2col
    v = torch.randn(48, generator=r)
    ops.rows_fill(X[:, 352:400], cu(v))
    close(X[:, 352:400], v[None].expand(Np, 48), 0)
    xin = torch.randn((Np, 16), generator=r)
    want = torch.zeros((Np, 144))
    fake_ops.im2col3x3(xin, want, Fc, h, w)
    got = rows_buf(Np, 144)
    XI = rows_buf(Np, 16); XI.copy_(xin)
    ops.im2col3x3(XI, got, Fc, h, w)
    close(got, want, 0)


# ---------------------------------------------------------------- tcgen05 GEMM (fp16 tier)
def _aligned(t, dtype=torch.float16):
    n, c = t.shape
    buf = torch.zeros((n, (c + 7) // 8 * 8 + 8), dtype=dtype, device="cuda")
    buf[:, :c] = t.to(dtype)
    return buf[:, :c]


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 128, 324), (1000, 486, 324), (257, 4, 192), (4224, 172, 600),
                                   (129, 600, 400), (64, 48, 8), (2000, 1152, 256), (700, 700, 128), (128, 2050, 128)])
def test_gemm_tensor_core(ops, M, N, K):
    r = torch.Generator().manual_seed(M * 7 + N * 3 + K)
    a = torch.randn((M, K), generator=r).half()
    w = (torch.randn((N, K), generator=r) / K ** 0.5).half()
    bias, dww, dwb = torch.randn(N, generator=r), torch.randn(N, generator=r) * 0.3, torch.randn(N, generator=r) * 0.3
    res = torch.randn((M, N), generator=r).half()
    A, Wt, R = _aligned(a), _aligned(w), _aligned(res)
    for kw in (dict(), dict(bias=bias, act="gelu"), dict(bias=bias, res=res, act="gelu", dw=(dww, dwb)),
               dict(bias=bias, act="relu", alpha=100.0), dict(res=res, alpha=0.5)):
        want = torch.zeros((M, N), dtype=torch.float32)
        fake_ops.gemm(a, w, want, **kw)
        kwc = {k: (tuple(cu(x) for x in v) if isinstance(v, tuple) else v) for k, v in kw.items()}
        if "bias" in kwc:
            kwc["bias"] = cu(bias)
        if "res" in kwc:
            kwc["res"] = R
        for odt in (torch.float16, torch.float32):
            got = _aligned(torch.zeros((M, N)), odt)
            ops.gemm(A, Wt, got, require_tc=True, **kwc)
            close(got, want, 2e-3, 2e-3)


@pytest.mark.parametrize("pair,cs", [(0, 1), (0, 2), (0, 3), (0, 4), (0, 5), (0, 8), (1, 2), (1, 4), (1, 6), (1, 8)])
@pytest.mark.parametrize("M,N,K", [(1408, 486, 324), (1300, 324, 486), (2176, 768, 512), (640, 128, 128), (1100, 600, 400), (900, 172, 600),
                                   (777, 193, 100), (512, 20, 64)])
def test_gemm_weight_stationary_cluster_sizes(ops, monkeypatch, pair, cs, M, N, K):
    """gemm_ws.cu with the mode (single CTAs / tcgen05 CTA pairs) and the cluster size pinned: multicast of A, the pair's
    relay and remote accumulator hand-back, TMA residual load / in-place update / TMA store, ragged last row tile (pair
    mode: a whole CTA of the last tile is out of range), column tiles that end inside a 32-column box, views into wider buffers."""
    monkeypatch.setenv("VFML_WS_CS", str(cs))
    monkeypatch.setenv("VFML_WS_PAIR", str(pair))
    r = torch.Generator().manual_seed(M + 3 * N + 7 * K + cs)
    a = torch.randn((M, K), generator=r).half()
    w = (torch.randn((N, K), generator=r) / K ** 0.5).half()
    bias, dww, dwb = torch.randn(N, generator=r), torch.randn(N, generator=r) * 0.3, torch.randn(N, generator=r) * 0.3
    res = torch.randn((M, N), generator=r).half()
    A, Wt = _aligned(a), _aligned(w)
    A.as_strided((M, A.stride(0) - K), (A.stride(0), 1), A.storage_offset() + K).fill_(50.0)       # poisoned row padding: the K tail
    Wt.as_strided((N, Wt.stride(0) - K), (Wt.stride(0), 1), Wt.storage_offset() + K).fill_(50.0)   # must come from TMA's zero fill
    wide = torch.full((M, 2 * ((N + 7) // 8 * 8) + 24), 7.0, dtype=torch.float16, device="cuda")
    R = wide[:, 8:8 + N]
    R.copy_(res)
    for kw in (dict(), dict(bias=bias), dict(bias=bias, act="gelu"), dict(bias=bias, res=res), dict(bias=bias, res=res, act="gelu"),
               dict(bias=bias, res=res, act="gelu", dw=(dww, dwb))):
        want = torch.zeros((M, N), dtype=torch.float32)
        fake_ops.gemm(a, w, want, **kw)
        kwc = {k: (tuple(cu(x) for x in v) if isinstance(v, tuple) else v) for k, v in kw.items()}
        if "bias" in kwc:
            kwc["bias"] = cu(bias)
        if "res" in kwc:
            kwc["res"] = R
        off = 8 + (N + 7) // 8 * 8 + 8
        got = wide[:, off:off + N]
        got.fill_(-3.0)
        ops.gemm(A, Wt, got, require_tc=True, **kwc)
        close(got, want, 2e-3, 2e-3)
        close(R, res, 0)                                          # the residual is untouched
        assert float(wide[:, off + N:].float().sub(7.0).abs().max()) == 0 and float(wide[:, 8 + N:off].float().sub(7.0).abs().max()) == 0


@pytest.mark.parametrize("M,N,K", [(1000, 172, 600), (4224, 486, 324), (7040, 128, 7040), (640, 20, 17)])
def test_gemm_fp32_split_tensor_cores(ops, M, N, K):
    """gemm_split.cu: fp32 product as three fp16 tensor-core products (hi / lo split, power-of-two operand scaling, fp32
    accumulation) + the fp32 epilogue pass; against the CPU double result: 22-bit operands -> <= 2e-6 of sum |a||w| (1e-5 at K = 7040),
    also with operands ~1e-4 (softmax rows) and ~1e3 that a plain fp16 cast would flush or overflow."""
    r = torch.Generator().manual_seed(M + N + K)
    for a_scale, w_scale in ((1.0, 1.0), (1e-4, 3.0), (2e3, 1e-2)):
        a = torch.randn((M, K), generator=r) * a_scale
        w = torch.randn((N, K), generator=r) * w_scale / K ** 0.5
        bias, dww, dwb = torch.randn(N, generator=r), torch.randn(N, generator=r) * 0.3, torch.randn(N, generator=r) * 0.3
        res = torch.randn((M, N), generator=r)
        ops.set_fp32_split(True)
        try:
            for kw in (dict(), dict(bias=bias, act="gelu"), dict(bias=bias, res=res, act="gelu", dw=(dww, dwb)), dict(res=res, alpha=0.5),
                       dict(bias=bias, act="relu", weights=True)):
                want = kw.get("alpha", 1.0) * (a.double() @ w.double().t())
                if "bias" in kw:
                    want = want + bias.double()
                if "res" in kw:
                    want = want + res.double()
                if kw.get("act") == "gelu":
                    want = torch.nn.functional.gelu(want)
                elif kw.get("act") == "relu":
                    want = torch.relu(want)
                if "dw" in kw:
                    want = torch.nn.functional.gelu(want + (dww.double() * want + dwb.double()))
                kwc = {k: (tuple(cu(x) for x in v) if isinstance(v, tuple) else cu(v) if torch.is_tensor(v) else v) for k, v in kw.items()}
                got = rows_buf(M, N)
                A = rows_buf(M, K, pad=16); A.copy_(a)
                ops.gemm(A, cu(w), got, **kwc)
                scale = float((a.double().abs() @ w.double().abs().t()).max()) + float(want.abs().max())
                err = float((got.double().cpu() - want).abs().max())
                # (the tensor core's fp32 accumulation is not round-to-nearest: over 3 x 7040 / 16 accumulation steps it shows)
                assert err <= (2e-6 if K <= 1024 else 1e-5) * scale, (kw.keys(), a_scale, err, scale)
        finally:
            ops.set_fp32_split(False)


def test_gemm_tensor_core_large_k_and_strided_views(ops):
    """attention-shaped products: K = 7040 (110 k-blocks through the stage ring), operands that are column
    slices of wider buffers."""
    r = torch.Generator().manual_seed(9)
    P = 1096
    attn = torch.softmax(torch.randn((P, P), generator=r) * 2, -1).half()
    vT = torch.randn((128, P), generator=r).half()
    mf = torch.randn((P, 128), generator=r).half()
    wide = torch.zeros((P, 512), dtype=torch.float16, device="cuda")
    wide[:, 256:384] = mf.cuda()
    want = torch.zeros((P, 128))
    fake_ops.gemm(attn, vT, want, res=mf, alpha=0.5)
    ops.gemm(_aligned(attn), _aligned(vT), wide[:, 384:512], res=wide[:, 256:384], alpha=0.5, require_tc=True)
    close(wide[:, 384:512], want, 3e-3, 2e-3)
    close(wide[:, 256:384], mf, 0)


def test_lookup_pair_matches_single(ops):
    r = np.random.default_rng(11)
    n, h, w = 3, 24, 40
    f1, f2, f3 = (r.standard_normal((n, 32, h, w)).astype(np.float32) for _ in range(3))
    fp, bp = ops.corr_volume_pyramid(cu(f1), cu(f2), 4), ops.corr_volume_pyramid(cu(f1), cu(f3), 4)
    cf = (omof.pixel_grid(n, h, w) + T((6 * r.standard_normal((n, 2, h, w))).astype(np.float32))).cuda()
    cb = (omof.pixel_grid(n, h, w) + T((6 * r.standard_normal((n, 2, h, w))).astype(np.float32))).cuda()
    both = rows_buf(2 * n * h * w, 324)
    ops.corr_lookup_pair(fp, bp, cf, cb, both[:n * h * w], both[n * h * w:])
    want_f = omof.corr_lookup(omof.corr_pyramid(T(f1), T(f2)), cf.cpu())
    want_b = omof.corr_lookup(omof.corr_pyramid(T(f1), T(f3)), cb.cpu())
    close(both[:n * h * w].view(n, h, w, 324).permute(0, 3, 1, 2), want_f, 5e-5)
    close(both[n * h * w:].view(n, h, w, 324).permute(0, 3, 1, 2), want_b, 5e-5)


@pytest.mark.parametrize("n,C,h,w", [(2, 256, 16, 24), (3, 64, 20, 31), (1, 256, 55, 128)])
def test_volume_fp16_tensor_core(ops, n, C, h, w):
    """fp16 tier: volume through the tcgen05 GEMM (fp16 operands, fp32 accumulate, fp16 storage) + pyramid."""
    r = np.random.default_rng(21)
    f1 = r.standard_normal((n, C, h, w)).astype(np.float32)
    f2 = r.standard_normal((n, C, h, w)).astype(np.float32)
    want = omof.corr_pyramid(T(f1).half().float(), T(f2).half().float())
    got = ops.corr_volume_pyramid(cu(f1), cu(f2), 4, torch.float16)
    for a, b in zip(got, want):
        assert a.dtype == torch.float16 and a.shape == b.shape
        close(a, b, 4e-3, 2e-3)


def test_corrblock_dropin(ops, golden):
    from video_flow_ml_b200.corr import CorrBlock
    g = golden("corr_lookup.npz")
    f1, f2, coords = cases.lookup_case()
    blk = CorrBlock(cu(f1), cu(f2), num_levels=4, radius=4)
    assert len(blk.corr_pyramid) == 4 and blk.corr_pyramid[0].shape == (768, 1, 16, 24)
    close(blk(cu(coords)), T(g["lookup"]), 5e-5)
    close(CorrBlock.corr(cu(f1), cu(f2))[0, :2], T(g["vol0"]), 2e-5)


# ---------------------------------------------------------------- tiled (gather-friendly) pyramid + lookup
def _tiled_case(ops, n, h, w, seed, spread=6.0):
    r = np.random.default_rng(seed)
    f1 = r.standard_normal((n, 64, h, w)).astype(np.float32)
    f2 = r.standard_normal((n, 64, h, w)).astype(np.float32)
    coords = omof.pixel_grid(n, h, w) + T((spread * r.standard_normal((n, 2, h, w))).astype(np.float32))
    # edge cases of the reference sampler: far out of range on both sides, exact integers, NaN
    flat = coords.view(n, 2, -1)
    flat[0, 0, 0], flat[0, 1, 1] = -1.0e9, 1.0e9
    flat[0, :, 2] = 3.0
    flat[0, 0, 3] = float("nan")
    flat[0, :, 4] = torch.tensor([w - 1.0, h - 1.0])
    flat[0, :, 5] = torch.tensor([-4.5, -3.25])
    flat[0, :, 6] = torch.tensor([w + 3.75, h + 2.5])
    return f1, f2, coords


@pytest.mark.parametrize("n,h,w", [(1, 16, 24), (3, 20, 31), (2, 47, 39), (1, 16, 16)])
@pytest.mark.parametrize("math_f32", [True, False])
def test_tiled_lookup_matches_oracle(ops, n, h, w, math_f32):
    f1, f2, coords = _tiled_case(ops, n, h, w, 11)
    pyr16 = ops.corr_volume_pyramid(cu(f1), cu(f2), 4, torch.float16)                 # row-major fp16 pyramid
    tp = ops.corr_tile_pyramid(pyr16[0].view(n * h * w, h * w), h, w)
    want = torch.zeros((n * h * w, 324), dtype=torch.float32, device="cuda")
    c = cu(coords)
    ops.corr_lookup(pyr16, c, 4, out=want)                                            # fp32 interpolation of the same fp16 texels
    got = rows_buf(n * h * w, 324, torch.float16, pad=4)
    ops.corr_lookup_tiled(tp, None, c, None, got, None, math_f32=math_f32)
    perm = ops.corr_lookup_perm("cuda")
    unperm = torch.empty_like(want)
    unperm[:, perm] = got.float()
    scale = want.abs().max().item()
    # fp32 math: one rounding to fp16 (2^-11 relative); fp16x2 math: three roundings
    tol = (1.0 if math_f32 else 3.0) * 2.0 ** -11 * scale + 1e-6
    err = (unperm - want).abs().max().item()
    assert err <= tol, f"max err {err:.3e} > {tol:.3e}"
    # and against the CPU oracle on fp32 volumes (fp16 storage error of the volume included)
    want_o = omof.corr_lookup(omof.corr_pyramid(T(f1), T(f2)), coords.nan_to_num(nan=-1.0e6))
    close(unperm.view(n, h, w, 324).permute(0, 3, 1, 2), want_o, 4.0 * 2.0 ** -11 * scale + 1e-6)


def test_tiled_lookup_full_size_integer_coordinates(ops):
    """BASELINE.json full size (436x1024 -> 55x128 features, one frame: a 99 MB level-0 volume).  Size-independent
    properties of the lookup (core/corr.py:31-56): at coordinates that are multiples of 8 every level's window centre is
    an integer, all bilinear weights are exactly 0 or 1, so (a) the tiled fp16 kernel, the rows kernel and a plain gather
    agree BIT-EXACTLY, and (b) the 81 taps of a level sum to the 9x9 box sum of that level's zero-padded map
    (a checksum that does not depend on the tap order)."""
    import torch.nn.functional as F
    n, h, w = 1, 55, 128
    P = h * w
    r = np.random.default_rng(123)
    f1 = r.standard_normal((n, 64, h, w)).astype(np.float32)
    f2 = r.standard_normal((n, 64, h, w)).astype(np.float32)
    cx = torch.from_numpy(8 * r.integers(0, w // 8, size=(n, 1, h, w))).float()
    cy = torch.from_numpy(8 * r.integers(0, (h + 7) // 8, size=(n, 1, h, w))).float()
    coords = cu(torch.cat([cx, cy], 1))
    pyr16 = ops.corr_volume_pyramid(cu(f1), cu(f2), 4, torch.float16)
    tp = ops.corr_tile_pyramid(pyr16[0].view(P, P), h, w)
    want = torch.zeros((P, 324), dtype=torch.float32, device="cuda")
    ops.corr_lookup(pyr16, coords, 4, out=want)
    perm = ops.corr_lookup_perm("cuda")
    for math_f32 in (False, True):
        got = rows_buf(P, 324, torch.float16, pad=4)
        ops.corr_lookup_tiled(tp, None, coords, None, got, None, math_f32=math_f32)
        unperm = torch.empty_like(want)
        unperm[:, perm] = got.float()
        assert torch.equal(unperm, want), (unperm - want).abs().max().item()
    ix, iy = coords[0, 0].reshape(-1).long(), coords[0, 1].reshape(-1).long()
    q = torch.arange(P, device="cuda")
    for l, lev in enumerate(pyr16):
        box = F.avg_pool2d(F.pad(lev.double(), (8, 8, 8, 8)), 9, stride=1, divisor_override=1)   # [P, 1, hl + 8, wl + 8]
        s_want = box[q, 0, (iy >> l) + 4, (ix >> l) + 4]
        s_got = unperm[:, 81 * l:81 * (l + 1)].double().sum(-1)
        assert torch.allclose(s_got, s_want, rtol=0, atol=1e-9), (l, (s_got - s_want).abs().max().item())


def test_tiled_volume_pair_launch(ops):
    """fmaps -> tiled pyramid in one call; both directions in one launch equal two single launches."""
    n, h, w = 2, 24, 40
    f1, f2, ca = _tiled_case(ops, n, h, w, 3)
    _, f3, cb = _tiled_case(ops, n, h, w, 4)
    ta = ops.corr_volume_pyramid_tiled(cu(f1), cu(f2))
    tb = ops.corr_volume_pyramid_tiled(cu(f1), cu(f3))
    pyr16 = ops.corr_volume_pyramid(cu(f1), cu(f2), 4, torch.float16)
    ref = ops.corr_tile_pyramid(pyr16[0].view(n * h * w, h * w), h, w)
    for a, b in zip(ta.levels, ref.levels):
        assert torch.equal(a, b)
    oa, ob = rows_buf(n * h * w, 324, torch.float16, 4), rows_buf(n * h * w, 324, torch.float16, 4)
    ops.corr_lookup_tiled(ta, tb, cu(ca), cu(cb), oa, ob)
    sa, sb = rows_buf(n * h * w, 324, torch.float16, 4), rows_buf(n * h * w, 324, torch.float16, 4)
    ops.corr_lookup_tiled(ta, None, cu(ca), None, sa, None)
    ops.corr_lookup_tiled(tb, None, cu(cb), None, sb, None)
    assert torch.equal(oa, sa) and torch.equal(ob, sb)


@pytest.mark.parametrize("n,C,h,w", [(2, 256, 55, 128), (1, 256, 47, 156), (3, 64, 20, 31), (1, 128, 8, 8), (2, 192, 9, 33),
                                     (1, 256, 17, 135), (6, 128, 24, 40), (1, 256, 8, 8)])
def test_corr_volume_fused_epilogue_is_bit_identical(ops, monkeypatch, n, C, h, w):
    """North star (1): the pyramid comes out of the volume GEMM's epilogue (corr_fused.cu: levels 0 and 1 from the TMEM
    accumulators, 2 and 3 from the tiled level 1).  Every byte of every level -- padding texels of the edge tiles included,
    the buffers start poisoned -- equals the unfused path (GEMM -> row-major fp16 level 0 -> corr_tile_pyramid_kernel), and
    level 0 matches the fp64 product.  Shapes: C1, C3 (C = 256: the engine's feature maps), odd sizes with M / N tails, the
    8 x 8 minimum, six pairs in one launch, every supported channel count (64, 128, 192, 256)."""
    r = np.random.default_rng(n * 1000 + h * w + C)
    f1 = cu(T(r.standard_normal((n, C, h, w)).astype(np.float32)))
    f2 = cu(T(r.standard_normal((n, C, h, w)).astype(np.float32)))

    def run(fused):
        monkeypatch.setenv("VFML_CORR_FUSED", "1" if fused else "0")
        out = ops.TiledPyramid([torch.full_like(l, 0xAB) for l in ops._tiled_levels(n * h * w, h, w, f1.device)], n, h, w)
        launches = ops._lib.launch_count()
        ops.corr_volume_pyramid_tiled(f1, f2, out=out)
        torch.cuda.synchronize()
        return out, ops._lib.launch_count() - launches

    fused, n_fused = run(True)
    plain, n_plain = run(False)
    assert n_fused == 4 and n_plain >= 4          # 2 x rows copies + (fused GEMM, levels 2-3) | (GEMM, pyramid)
    for l, (a, b) in enumerate(zip(fused.levels, plain.levels)):
        assert torch.equal(a, b), (l, int((a != b).sum()))
    # level 0 against the fp64 product of the fp16-rounded features (texel (x, y) of tile (x // 4, y // 4), x-major in the tile)
    a16, b16 = f1.half().double().flatten(2), f2.half().double().flatten(2)
    want = torch.einsum("ncp,ncq->npq", a16, b16) / np.sqrt(C)
    nty = (h + 3) // 4
    lv0 = fused.levels[0].view(torch.float16).view(n, h * w, -1, 16)
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    tile = ((xs // 4) * nty + ys // 4).flatten().to(f1.device)
    elem = ((xs % 4) * 4 + ys % 4).flatten().to(f1.device)
    got = lv0[:, :, tile, elem].double()
    assert (got - want).abs().max().item() < 2e-2 * max(1.0, want.abs().max().item() / 16)


# ---------------------------------------------------------------- tensor-core depthwise convolution
@pytest.mark.parametrize("k,C,n,h,w", [(15, 324, 2, 55, 128), (15, 128, 1, 20, 31), (7, 512, 1, 47, 39), (7, 64, 3, 70, 33), (15, 12, 1, 8, 8)])
def test_dwconv_tc_matches_fp64(ops, k, C, n, h, w):
    r = np.random.default_rng(k + C)
    x = T(r.standard_normal((n * h * w, C)).astype(np.float32))
    wt = T((r.standard_normal((k * k, C)) / k).astype(np.float32))
    b = T(r.standard_normal(C).astype(np.float32))
    xb, yb = rows_buf(n * h * w, C, torch.float16, pad=(-C) % 8 + 8), rows_buf(n * h * w, C, torch.float16, pad=(-C) % 8 + 8)
    xb.copy_(x.cuda())
    packed = ops.dwconv_pack_weights(wt.cuda(), k)
    ops.dwconv_tc_res_gelu(xb, yb, n, h, w, packed, b.cuda(), k)
    # reference in float64 on the fp16-rounded operands the kernel sees
    xi = xb.cpu().double().reshape(n, h, w, C).permute(0, 3, 1, 2)
    w64 = wt.half().double().t().reshape(C, 1, k, k)
    conv = torch.nn.functional.conv2d(xi, w64, b.double(), padding=k // 2, groups=C)
    want = torch.nn.functional.gelu(xi + conv).permute(0, 2, 3, 1).reshape(n * h * w, C)
    err = (yb.cpu().double() - want).abs().max().item()
    assert err <= 2.0 ** -10 * max(1.0, want.abs().max().item()), err      # one fp16 rounding of the result
    # the padding columns of the rows buffer stay zero
    base = yb._base if yb._base is not None else yb
    assert base[:, C:].abs().max().item() == 0
    # and it agrees with the FFMA kernel
    y2 = rows_buf(n * h * w, C, torch.float16, pad=(-C) % 8 + 8)
    ops.dwconv_res_gelu(xb, y2, n, h, w, wt.cuda(), b.cuda(), k)
    assert (y2.float() - yb.float()).abs().max().item() <= 1e-2 * max(1.0, want.abs().max().item())


# ---------------------------------------------------------------- volume-free lookup (alt_cuda_corr)
@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-4), (torch.float16, 2e-2)])
@pytest.mark.parametrize("B,N,H1,W1,H2,W2,C,r", [(2, 1, 16, 24, 16, 24, 256, 4), (1, 2, 9, 7, 5, 11, 128, 3),
                                                  (1, 1, 13, 5, 13, 5, 512, 4), (3, 1, 6, 6, 3, 3, 256, 1)])
def test_alt_corr_forward_matches_oracle(ops, dt, tol, B, N, H1, W1, H2, W2, C, r):
    g = np.random.default_rng(B * 1000 + C + r)
    f1 = T(g.standard_normal((B, H1, W1, C)).astype(np.float32)).to(dt)
    f2 = T(g.standard_normal((B, H2, W2, C)).astype(np.float32)).to(dt)
    coords = T((g.uniform(-6, 6, (B, N, H1, W1, 2)) + np.array([W2, H2]) * g.uniform(0, 1, (B, N, H1, W1, 2))).astype(np.float32))
    coords[0, 0, 0, 0] = torch.tensor([2.0, 3.0])               # exact integers: dx = dy = 0
    coords[0, 0, 0, 1] = torch.tensor([-1.0e9, 4.0])
    coords[0, 0, 0, 2] = torch.tensor([W2 - 1.0, H2 - 1.0])
    want = omof.alt_corr_forward(f1.float(), f2.float(), coords, r)
    got, = ops.alt_corr_forward(cu(f1), cu(f2), cu(coords), r)
    assert got.shape == (B, N, (2 * r + 1) ** 2, H1, W1) and got.dtype == torch.float32
    close(got, want, tol * float(C) ** 0.5)
    nanc = coords.clone()
    nanc[0, 0, 0, 3, 0] = float("nan")                          # UB in the reference; here: zeros, nothing else disturbed
    got2, = ops.alt_corr_forward(cu(f1), cu(f2), cu(nanc), r)
    assert torch.all(got2[0, 0, :, 0, 3] == 0) and torch.isfinite(got2).all()
    got2[0, 0, :, 0, 3] = got[0, 0, :, 0, 3]
    assert torch.equal(got2, got)


def test_alternate_corrblock_dropin(ops, golden):
    from video_flow_ml_b200.corr import AlternateCorrBlock
    g = golden("corr_lookup.npz")
    f1, f2, coords = cases.lookup_case()
    blk = AlternateCorrBlock(cu(f1), cu(f2), num_levels=4, radius=4)
    out = blk(cu(coords))
    assert out.shape == (2, 324, 16, 24) and out.dtype == torch.float32
    close(out, T(g["lookup"]), 5e-5)


def test_alt_corr_errors(ops):
    f = torch.zeros((1, 4, 4, 96), device="cuda")
    c = torch.zeros((1, 1, 4, 4, 2), device="cuda")
    with pytest.raises(RuntimeError, match="unsupported"):
        ops.alt_corr_forward(f, f, c, 4)
    with pytest.raises(ValueError):
        ops.alt_corr_forward(f.permute(0, 2, 1, 3), f, c, 4)
    with pytest.raises(RuntimeError, match="radius"):
        ops.alt_corr_forward(torch.zeros((1, 4, 4, 128), device="cuda"), torch.zeros((1, 4, 4, 128), device="cuda"), c, 5)


# ---------------------------------------------------------------- split-K aggregate GEMM
@pytest.mark.parametrize("M,K,N,S", [(7040, 7040, 128, 5), (1000, 1536, 128, 3), (384, 768, 128, 2)])
def test_gemm_splitk_matches_single_pass(ops, M, K, N, S):
    g = torch.Generator(device="cuda").manual_seed(M + K)
    a = torch.softmax(4 * torch.randn((M, K), device="cuda", generator=g), -1).half()       # an affinity matrix
    buf = torch.zeros((N, K + 8), device="cuda", dtype=torch.float16)
    w = buf[:, :K]
    w.copy_(torch.randn((N, K), device="cuda", generator=g).half())
    res = rows_buf(M, N, torch.float16)
    res.copy_(torch.randn((M, N), device="cuda", generator=g).half())
    want = res.double() + 0.5 * (a.double() @ w.double().t())
    out = rows_buf(M, 2 * N, torch.float16)[:, N:]                  # a column range of a wider buffer
    part = torch.empty(S * M * N, device="cuda", dtype=torch.float32)
    ops.gemm_splitk(a, w, out, S, part, res=res, alpha=0.5)
    close(out, want, 2e-3, 1e-3)
    one = rows_buf(M, N, torch.float16)
    ops.gemm(a, w, one, res=res, alpha=0.5)
    close(out, one, 2e-3, 1e-3)
    with pytest.raises(ValueError):
        ops.gemm_splitk(a, w, out, 7, part, res=res)


def test_engine_split_factor():
    from video_flow_ml_b200.mof import MOFEngine
    assert MOFEngine._splits(7040, 7040) == 2 and MOFEngine._splits(7040, 14080) == 2
    assert MOFEngine._splits(384, 384) == 1 and MOFEngine._splits(21600, 21600) == 1 and MOFEngine._splits(7040, 7000) == 1


# ---------------------------------------------------------------- LayerNorm (Twins blocks)
@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-6), (torch.float16, 2e-3)])
@pytest.mark.parametrize("shape", [(5, 1000, 128), (3, 77, 256), (1, 9, 64), (2, 33, 320), (1, 5, 512)])
def test_layer_norm_rows_matches_torch(ops, dt, tol, shape):
    g = torch.Generator(device="cuda").manual_seed(shape[1])
    x = (3 * torch.randn(shape, device="cuda", generator=g) + 1.5).to(dt)
    w = (1 + 0.2 * torch.randn(shape[-1], device="cuda", generator=g)).to(dt)
    b = (0.3 * torch.randn(shape[-1], device="cuda", generator=g)).to(dt)
    want = torch.nn.functional.layer_norm(x.double(), (shape[-1],), w.double(), b.double(), 1e-6)
    got = ops.layer_norm_rows(x, w, b, 1e-6)
    assert got.shape == x.shape and got.dtype == dt
    close(got, want, tol * 4, tol)
    with pytest.raises(ValueError):
        ops.layer_norm_rows(x.transpose(1, 2), w, b, 1e-6)
    with pytest.raises(RuntimeError):
        ops.layer_norm_rows(torch.zeros((4, 10), device="cuda", dtype=dt), torch.ones(10, device="cuda", dtype=dt),
                            torch.zeros(10, device="cuda", dtype=dt), 1e-6)


def test_peg3x3_matches_conv2d():
    """Twins PEG kernel (x + dwconv3x3(x) + b, token-major) against torch's conv2d in fp64 on the CPU."""
    from video_flow_ml_b200 import ops
    g = torch.Generator().manual_seed(3)
    for (B, H, W, C) in ((1, 20, 20, 256), (2, 13, 7, 128), (3, 55, 64, 8)):
        x = torch.randn((B, H * W, C), generator=g)
        wt = torch.randn((C, 1, 3, 3), generator=g) * 0.3
        b = torch.randn((C,), generator=g)
        m = x.double().transpose(1, 2).reshape(B, C, H, W)
        want = (m + torch.nn.functional.conv2d(m, wt.double(), b.double(), padding=1, groups=C)).flatten(2).transpose(1, 2)
        for dt, tol in ((torch.float32, 1e-5), (torch.float16, 2e-2)):
            got = ops.peg3x3(x.cuda().to(dt).contiguous(), B, H, W, wt.reshape(C, 9).t().contiguous().cuda().to(dt), b.cuda().to(dt))
            ref = want if dt == torch.float32 else (x.half().double().transpose(1, 2).reshape(B, C, H, W) + torch.nn.functional.conv2d(
                x.half().double().transpose(1, 2).reshape(B, C, H, W), wt.half().double(), b.half().double(), padding=1, groups=C)).flatten(2).transpose(1, 2)
            assert (got.double().cpu() - ref).abs().max() <= tol, (B, H, W, C, dt)


def test_twins_is_batch_independent():
    """The encoders are per-image functions (the per-frame feature cache relies on it): one frame alone gives the same
    features as inside a batch, in both tiers, also for small tiles.  (Regression: a library depthwise kernel returned
    garbage for batch 1, 20x20, 256 channels, fp16.)"""
    from oracle import mof as omof
    from video_flow_ml_b200.twins import TwinsStage01
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    g = torch.Generator().manual_seed(5)
    for (H, W) in ((160, 160), (128, 128), (288, 448)):
        x = (torch.rand((5, 3, H, W), generator=g) * 2 - 1).cuda()
        for dt, tol in ((torch.float32, 1e-4), (torch.float16, 5e-2)):
            for prefix in ("fnet.svt", "cnet.svt"):
                enc = TwinsStage01(sd, prefix, "cuda", dt)
                full = enc(x)
                assert torch.isfinite(full).all()
                for b in (1, 2):
                    assert (enc(x[:b]) - full[:b]).abs().max() <= tol, (H, W, dt, prefix, b)


@pytest.mark.parametrize("B,Nq,Nk,heads", [(3, 49, 49, 4), (2, 7040, 416, 8), (1, 1600, 25, 4), (2, 100, 130, 8), (1, 64, 64, 1)])
def test_attention_hd32_matches_fp64(B, Nq, Nk, heads):
    """Fused attention kernel (Twins LSA / GSA shapes, ragged key counts, slices of fused projections used in place)."""
    from video_flow_ml_b200 import ops
    g = torch.Generator().manual_seed(Nq + Nk)
    C = heads * 32
    qkv = (torch.randn((B, Nq, 3 * C), generator=g) * 1.5).half().cuda()
    kv = (torch.randn((B, Nk, 2 * C), generator=g) * 1.5).half().cuda()
    q, k, v = qkv[..., C:2 * C], kv[..., 0:C], kv[..., C:2 * C]
    got = ops.attention_hd32(q, k, v, heads)
    qd, kd, vd = (t.double().cpu().view(B, -1, heads, 32).transpose(1, 2) for t in (q, k, v))
    want = (torch.softmax(qd @ kd.transpose(-1, -2) * 32 ** -0.5, -1) @ vd).transpose(1, 2).reshape(B, Nq, C)
    assert got.shape == (B, Nq, C) and got.dtype == torch.float16
    err = (got.double().cpu() - want).abs().max().item()
    assert err <= 4e-3, err
    with pytest.raises(ValueError):
        ops.attention_hd32(q.float(), k, v, heads)
