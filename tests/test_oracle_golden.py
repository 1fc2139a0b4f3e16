"""The oracle restatements against golden vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

import cases
from oracle import mof, post, twins

T = torch.from_numpy
torch.set_grad_enabled(False)


def close(a, b, atol, rtol=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    err = np.abs(a - b)
    assert err.max() <= atol + rtol * np.abs(b).max(), f"max err {err.max():.3e} (ref max {np.abs(b).max():.3e})"


def test_corr_volume_pyramid_lookup(golden):
    g = golden("corr_lookup.npz")
    f1, f2, coords = cases.lookup_case()
    vol = mof.corr_volume(T(f1), T(f2))
    close(vol[0, :2].numpy(), g["vol0"], 1e-5)
    pyr = mof.corr_pyramid(T(f1), T(f2))
    assert [tuple(p.shape[-2:]) for p in pyr] == [(16, 24), (8, 12), (4, 6), (2, 3)]
    close(pyr[3][:64].numpy(), g["pyr3"], 1e-5)
    out = mof.corr_lookup(pyr, T(coords))
    assert out.shape == (2, 324, 16, 24) and out.is_contiguous()
    close(out.numpy(), g["lookup"], 2e-5)


def test_lookup_tap_order_q13():
    """channel l*81+i*9+j samples (x+i-4, y+j-4): x offset is the slow index."""
    h, w = 16, 24
    f1, f2, _ = cases.lookup_case()
    pyr = mof.corr_pyramid(T(f1[:1]), T(f2[:1]))
    c = mof.pixel_grid(1, h, w)
    out = mof.corr_lookup(pyr, c)
    vol = pyr[0].view(h, w, h, w)
    y, x, i, j = 7, 9, 6, 1
    assert abs(out[0, i * 9 + j, y, x].item() - vol[y, x, y + j - 4, x + i - 4].item()) < 1e-5


def test_update_block(golden, mof_weights):
    g = golden("update_block.npz")
    c = {k: T(v) for k, v in cases.update_case().items()}
    Fc, _, h, w = c["net"].shape
    attn = mof.gma_attention(c["inp"], mof_weights)
    close(attn[:, 0, :8].numpy(), g["attn_rows"], 1e-6)
    c0 = mof.pixel_grid(Fc, h, w)
    close(mof.pcblock(c["fcorr"], mof_weights, "update_block.encoder.convc1", (1, 15)).numpy(), g["convc1"], 1e-4)
    net, mhs, mask, delta = mof.update_block(c["net"], c["mhs"], c["inp"], c["fcorr"], c["bcorr"], c["fflow"],
                                             c["bflow"], c0, attn, 1, mof_weights)
    close(net.numpy(), g["net"], 2e-4)
    close(mhs.numpy(), g["mhs"], 2e-4)
    close(delta.numpy(), g["delta"], 2e-4)
    close(mask[:, ::37].numpy(), g["mask"], 2e-2, 1e-5)
    net, mhs, mask, delta = mof.update_block(c["net"], None, c["inp"], c["fcorr"], c["bcorr"], c["fflow"],
                                             c["bflow"], c0, attn, 1, mof_weights)
    close(net.numpy(), g["net_first"], 2e-4)
    close(mhs.numpy(), g["mhs_first"], 2e-4)
    close(delta.numpy(), g["delta_first"], 2e-4)


def test_convex_upsample(golden):
    fl, mk = cases.upsample_case()
    close(mof.convex_upsample8(T(fl), T(mk)).numpy(), golden("upsample.npz")["up"], 1e-5)


def test_twins_restatement(golden, mof_weights):
    g = golden("twins.npz")
    x = T(cases.twins_case())
    close(twins.twins_svt_forward(x, mof_weights, "fnet.svt").numpy()[:, ::4], g["fnet"], 1e-4)
    close(twins.twins_svt_forward(x, mof_weights, "cnet.svt").numpy()[:, ::4], g["cnet"], 1e-4)


def test_mof_end_to_end_wrapper(golden, mof_weights):
    """Full wrapper stack: window choice, /255, pad, 32 iterations, unpad, index T-2 (Q2)."""
    g = golden("mof_e2e.npz")
    frames = cases.mof_case()
    win = mof.frames_to_window(frames, 2, 5)
    flows = mof.mof_forward(mof.pad_frames(win), mof_weights)
    close(flows.numpy()[0, :, :, ::2, ::2], g["all_flows_2"], 1e-2)
    f2 = mof.unpad(flows, 124, 188)[0, 3].permute(1, 2, 0).numpy()
    epe = np.sqrt(((f2 - g["flow_2"]) ** 2).sum(-1)).max()
    assert epe < 1e-2, epe
    f5 = mof.core_compute_flow(mof.frames_to_window(frames, 5, 5), mof_weights).permute(1, 2, 0).numpy()
    assert np.sqrt(((f5 - g["flow_5"]) ** 2).sum(-1)).max() < 1e-2
    u8 = cases.mof_case(as_float=False)
    f0 = mof.core_compute_flow(mof.frames_to_window(u8, 0, 5), mof_weights).permute(1, 2, 0).numpy()
    assert np.sqrt(((f0 - g["flow_u8_0"]) ** 2).sum(-1)).max() < 1e-2


def test_mof_end_to_end_wellconditioned(golden):
    """Second weight set (gain 1.0, make_golden_wellcond.py): the fixture the GPU tiers are held to."""
    g = golden("mof_e2e_g10.npz")
    sd = mof.mof_synthetic_weights(0, gain=1.0)
    f2 = mof.core_compute_flow(mof.frames_to_window(cases.mof_case(), 2, 5), sd).permute(1, 2, 0).numpy()
    assert np.sqrt(((f2 - g["flow_2"]) ** 2).sum(-1)).max() < 1e-3


def test_host_logic(golden):
    with open(os.path.join(os.path.dirname(cases.__file__), "host.json")) as f:
        host = json.load(f)
    for key, want in host["windows"].items():
        n, idx, Tn = map(int, key.split(","))
        assert mof.window_indices(n, idx, Tn) == want, key
    for key, want in host["tiles"].items():
        wd, ht = map(int, key.split("x"))
        assert list(mof.tile_grid(wd, ht)) == want, key
    assert mof.pad_amounts(436, 1024) == (0, 0, 2, 2)
    assert mof.pad_amounts(376, 1242) == (3, 3, 0, 0)


def test_encoders_bit_exact(golden):
    g = golden("post.npz")
    ff = cases.flow_field_case()
    h, w = ff.shape[:2]
    assert np.array_equal(post.encode_gamedev(ff, w, h), g["gamedev"])
    assert np.array_equal(post.encode_gamedev(ff, 3840, 2160), g["gamedev_4k"])
    assert np.array_equal(post.encode_mv_rg8(ff), g["rg8"])
    assert np.array_equal(post.encode_mv_rg8(ff, 16.0), g["rg8_c16"])
    assert np.array_equal(post.decode_mv_rg8(g["rg8"]), g["rg8_decoded"])


@pytest.mark.parametrize("name,kw", [("taa_bilateral", dict(use_flow=True, use_bilateral=True)),
                                     ("taa_bilinear", dict(use_flow=True, use_bilateral=False)),
                                     ("taa_simple", dict(use_flow=False, use_bilateral=False))])
def test_taa(golden, name, kw):
    g = golden("post.npz")
    frames, flows = cases.taa_case()
    t = post.TAAOracle(alpha=0.1)
    prev = None
    for i, f in enumerate(frames):
        o = t.apply(f, prev, alpha=0.1, seq="s", **kw)
        want = g[f"{name}_{i}"]
        assert o.dtype == want.dtype, (i, o.dtype, want.dtype)
        assert np.array_equal(o, want), (name, i, np.abs(o - want).max())
        prev = flows[i].copy()


def test_bof_oracle_matches_reference_golden(golden):
    """oracle/bof.py against the output of the unmodified reference BOFNet (tests/golden/make_golden_bof.py)."""
    from oracle import bof as obof
    g = golden("bof_e2e.npz")
    sd = obof.bof_synthetic_weights(0)
    frames = cases.mof_case()
    win = mof.frames_to_window(frames, 2, 3)
    flows = mof.unpad(obof.bof_forward(mof.pad_frames(win), sd), *win.shape[-2:])[0]
    assert np.abs(flows.numpy() - g["flows_2"]).max() < 1e-4
    assert np.abs(obof.core_compute_flow(win, sd).numpy() - g["flows_2"][1]).max() < 1e-4


def test_alt_corr_oracle_matches_reference_corrblock(golden):
    """The volume-free restatement (alt_cuda_corr + AlternateCorrBlock) against the reference CorrBlock output:
    the two reference blocks are documented as interchangeable (VideoFlow/README.md:112)."""
    g = golden("corr_lookup.npz")
    f1, f2, coords = cases.lookup_case()
    out = mof.alt_corr_block(T(f1), T(f2), T(coords))
    assert out.shape == (2, 324, 16, 24)
    close(out.numpy(), g["lookup"], 3e-5)


def test_memflow_oracle_matches_reference_golden(golden):
    """oracle/memflow.py against the reference MemFlowNet + InferenceCore (make_golden_memflow.py): a 3-step
    sequence (memory empty / one stored frame / replaced) and the application's fresh-core end=True call."""
    from oracle import memflow as omem
    g = golden("memflow_e2e.npz")
    sd = omem.memflow_synthetic_weights(0)
    frames = cases.mof_case()
    x = torch.stack([T(f).permute(2, 0, 1) for f in frames[:4]])[None]
    x = mof.pad_frames(2 * (x / 255.0) - 1.0)
    core = omem.InferenceCore(sd)
    for t in range(2):           # step 2 is covered on the GPU (CPU time)
        low, up = core.step(x[:, t:t + 2])
        close(low.numpy(), g[f"seq_low_{t}"], 2e-5)
        close(up.numpy(), g[f"seq_up_{t}"], 2e-4)
    assert core.mem_k.shape == (1, 128, 16 * 24) and core.mem_v.shape == (1, 128, 16 * 24)
    flow = omem.core_compute_flow(mof.frames_to_window(frames, 2, 3), sd)
    assert flow.shape == (2, 124, 188)
    close(flow.numpy(), g["core_flow"], 2e-4)


# ---------------------------------------------------------------------------- round 2 fixtures
def _post2_inputs():
    ff = cases.flow_field_case()
    fin = np.nan_to_num(cases.flow_field_case(seed=31, h=45, w=61), nan=0.5, posinf=3.0, neginf=-3.0)
    fe = np.nan_to_num(cases.flow_field_case(seed=37, h=64, w=96), nan=1.0, posinf=2.0, neginf=-2.0)
    frames, flows = cases.taa_case(seed=41, n=3, h=33, w=47)
    fl = flows[0] * 2
    fl[3, 3] = [100.0, 0.0]
    fl[4, 4] = [0.0, -100.0]
    return ff, fin, fe, frames, fl


def test_remaining_encoders_bit_exact(golden):
    """rgb8 ('rgb+') encode / decode, HSV, torchvision: oracle == unmodified reference, bit for bit (post2.npz)."""
    g = golden("post2.npz")
    ff, fin, _, _, _ = _post2_inputs()
    with np.errstate(all="ignore"):
        assert np.array_equal(post.encode_mv_rgb8(ff), g["rgb8"])
        assert np.array_equal(post.encode_mv_rgb8(ff, 8.0), g["rgb8_c8"])
        assert np.array_equal(post.decode_mv_rgb8(g["rgb8"]), g["rgb8_decoded"])
        assert np.array_equal(post.encode_mv_rgb8(ff, 64.0), g["mv_default"])          # encode_motion_vectors default = rgb8, clamp 64
        assert np.array_equal(post.decode_mv_rgb8(g["mv_default"], 64.0), g["mv_default_decoded"])
        assert np.array_equal(post.encode_hsv(ff), g["hsv"])
        assert np.array_equal(post.encode_hsv(fin), g["hsv_fin"])
        assert np.array_equal(post.encode_hsv(np.zeros((5, 7, 2), np.float32)), g["hsv_zero"])
        assert np.array_equal(post.encode_torchvision(fin), g["tv"])


def test_hsv2rgb_restatement_matches_opencv():
    """The third-party step of the HSV encoder: exhaustive (H, S) at V = 255 and random V, SIMD body and scalar tail."""
    cv2 = pytest.importorskip("cv2")
    hs = np.zeros((181, 256, 3), np.uint8)
    hs[..., 0] = np.arange(181)[:, None]
    hs[..., 1] = np.arange(256)[None]
    hs[..., 2] = 255
    assert np.array_equal(post.hsv2rgb_u8(hs), cv2.cvtColor(hs, cv2.COLOR_HSV2RGB))
    r = np.random.default_rng(0)
    for w in (1, 7, 31, 32, 33, 53, 100):
        x = r.integers(0, 256, (9, w, 3)).astype(np.uint8)
        x[..., 0] %= 181
        assert np.array_equal(post.hsv2rgb_u8(x), cv2.cvtColor(x, cv2.COLOR_HSV2RGB)), w


def test_flo_lod_quality_oracle(golden, tmp_path):
    g = golden("post2.npz")
    _, fin, fe, frames, fl = _post2_inputs()
    assert np.array_equal(np.frombuffer(post.flo_bytes(fin), np.uint8), g["flo_bytes"])
    for name, f in (("lod_odd", fin), ("lod_even", fe)):
        lods = post.lod_pyramid(f, 5)
        assert len(lods) == 5
        for i in range(1, 5):
            assert np.array_equal(lods[i], g[f"{name}_{i}"]), (name, i)
    assert np.array_equal(post.quality_frame(frames[0], frames[1], fl, 0.7), g["quality"])
    assert np.array_equal(post.quality_frame(frames[0], frames[0], np.zeros_like(fl), 0.7), g["quality_same"])
    small = np.ascontiguousarray(fl[::2, ::2][:16, :23])
    assert np.array_equal(post.quality_frame(frames[0], frames[1], small, 0.7), g["quality_resized"])
    # the product's file handler is host code: byte-identical .flo, lossless round trips, no GPU needed
    from video_flow_ml_b200.storage import FlowCacheManager, FlowFileHandler
    p = str(tmp_path / "a.flo")
    FlowFileHandler.save_flow_flo(fin, p)
    assert open(p, "rb").read() == g["flo_bytes"].tobytes()
    assert np.array_equal(FlowFileHandler.load_flow_flo(p), fin)
    open(p, "wb").write(b"XXXX" + b"\0" * 8)
    with pytest.raises(ValueError):
        FlowFileHandler.load_flow_flo(p)
    m = FlowCacheManager()
    d = str(tmp_path / "cache")
    assert m.check_cache_exists(d, 2) == (False, None, [0, 1])
    m.save_flow_to_cache(fin, d, 0, "both")
    assert m.check_cache_exists(d, 2) == (False, None, [1])
    m.save_flow_to_cache(torch.from_numpy(fe), d, 1, "npz")
    assert m.check_cache_exists(d, 2) == (True, "npz", [])
    assert np.array_equal(m.load_cached_flow(d, 0), fin) and np.array_equal(m.load_cached_flow(d, 0, "flo"), fin)
    z = FlowFileHandler.load_flow_npz(os.path.join(d, "flow_frame_000001.npz"))
    assert int(z["frame_idx"]) == 1 and tuple(z["shape"]) == fe.shape and np.array_equal(z["flow"], fe)
    with pytest.raises(FileNotFoundError):
        m.load_cached_flow(d, 5)
    with pytest.raises(ValueError):
        m.load_cached_flow(d, 0, "bmp")
    m.save_flow_lods(post.lod_pyramid(fin, 3), d, 0)
    assert np.array_equal(m.load_flow_lod(d, 0, 2), g["lod_odd_2"]) and not m.check_flow_lods_exist(d, 2, 3)


def test_large_motion_and_fullsize_fixtures(golden):
    """The drift weights (gain 1.0 + flow-head bias drift) against the unmodified reference wrapper: the oracle follows
    coordinates that travel ~20 low-res texels through the border (mof_e2e_drift.npz), tolerance = a few times the
    fixture's own 1-ulp sensitivity (stored in the fixture)."""
    g = golden("mof_e2e_drift.npz")
    assert float(g["ulp_sens"]) < 5e-3 and 0.5 < float(g["floor_fp16"]) < 5.0
    sd = mof.mof_synthetic_weights(0, gain=1.0, drift=mof.DRIFT)
    frames = cases.mof_case()
    win = mof.frames_to_window(frames, 2, 5)
    flow = mof.core_compute_flow(win, sd).permute(1, 2, 0).numpy()
    assert np.abs(g["flow_2"]).max() > 100.0                       # >= 12 low-res texels of motion
    assert np.sqrt(((flow - g["flow_2"]) ** 2).sum(-1)).max() <= 1e-2
