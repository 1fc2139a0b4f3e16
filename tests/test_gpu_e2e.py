"""End-to-end parity of the CUDA path (through the drop-in wrappers and the C ABI) against the golden
outputs of the unmodified reference wrapper stack (tests/golden/mof_e2e.npz).  pytest -m gpu."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import mof as omof

pytestmark = pytest.mark.gpu


def epe(a, b):
    return np.sqrt(((a - b) ** 2).sum(-1)).max()


@pytest.fixture(scope="module")
def processor(mof_weights):
    from video_flow_ml_b200.processing import VideoFlowProcessor
    p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp32")
    with pytest.raises(RuntimeError):
        p.compute_optical_flow([np.zeros((64, 64, 3), np.uint8)] * 5, 0)
    p.core.load_state_dict({"module." + k: v for k, v in mof_weights.items()})
    return p


@pytest.fixture(scope="module")
def processor_g10():
    """Well-conditioned weight set (tests/golden/make_golden_wellcond.py)."""
    from video_flow_ml_b200.processing import VideoFlowProcessor
    p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp32")
    p.core.load_state_dict(omof.mof_synthetic_weights(0, gain=1.0))
    return p


def test_mof_fp32_matches_reference_golden(processor_g10, golden):
    """32 iterations, fp32 tier, through the wrapper stack: max EPE delta <= 1e-2 px vs the output of the
    unmodified reference (north-star fp32 tolerance)."""
    g = golden("mof_e2e_g10.npz")
    frames = cases.mof_case()
    for idx in (2, 5):
        flow = processor_g10.compute_optical_flow(list(frames), idx)
        assert flow.shape == (124, 188, 2) and flow.dtype == np.float32
        assert epe(flow, g[f"flow_{idx}"]) <= 1e-2, (idx, epe(flow, g[f"flow_{idx}"]))
    win = processor_g10.prepare_frame_sequence(list(frames), 2)
    flows = processor_g10.core.model.forward(omof.pad_frames(win.cpu()).cuda())
    assert np.abs(flows.cpu().numpy()[0, :, :, ::2, ::2] - g["all_flows_2"]).max() <= 1e-2


def test_mof_fp32_stress_weights(processor, golden):
    """Ill-conditioned weight set (update-block gain 1.4): the reference's own output moves by 3e-3 px
    under a 1-ulp perturbation of the encoder outputs (make_golden_wellcond.py docstring), so the bound
    here is 5e-2 px; it still pins window choice, /255 branches, padding, unpadding and flow index."""
    g = golden("mof_e2e.npz")
    frames = cases.mof_case()
    for idx in (2, 5):
        flow = processor.compute_optical_flow(list(frames), idx)
        assert epe(flow, g[f"flow_{idx}"]) <= 5e-2, (idx, epe(flow, g[f"flow_{idx}"]))
    u8 = cases.mof_case(as_float=False)
    f0 = processor.compute_optical_flow(list(u8), 0)
    assert epe(f0, g["flow_u8_0"]) <= 5e-2


def test_mof_all_flows_match_reference_golden(processor, golden):
    g = golden("mof_e2e.npz")
    win = processor.prepare_frame_sequence(list(cases.mof_case()), 2)
    flows = processor.core.model.forward(omof.pad_frames(win.cpu()).cuda())
    assert flows.shape == (1, 6, 2, 128, 192)
    err = np.abs(flows.cpu().numpy()[0, :, :, ::2, ::2] - g["all_flows_2"]).max()
    assert err <= 5e-2, err


def test_core_validation_errors(processor):
    core = processor.core
    with pytest.raises(ValueError):
        core.compute_flow_from_tensor(np.zeros((1, 5, 3, 64, 64), np.float32))
    with pytest.raises(ValueError):
        core.compute_flow_from_tensor(torch.zeros((1, 5, 3, 64, 64)))             # CPU tensor
    with pytest.raises(ValueError):
        core.compute_flow_from_tensor(torch.zeros((5, 3, 64, 64), device="cuda"))
    with pytest.raises(ValueError):
        core.compute_flow_from_tensor(torch.zeros((2, 5, 3, 64, 64), device="cuda"))
    with pytest.raises(ValueError):
        core.compute_flow_from_tensor(torch.zeros((1, 5, 4, 64, 64), device="cuda"))
    assert core.get_model_info()["status"] == "loaded"


def test_tiled_equals_per_tile_clips(processor_g10):
    processor = processor_g10
    """Tile mode = independent clips per disjoint tile, pasted (reference :231-283).  Uses a small tile
    size through calculate_tile_grid's parameter by monkeypatching the default."""
    frames = cases.synthetic_clip(5, 192, 352, seed=5)
    orig = processor.calculate_tile_grid
    processor.calculate_tile_grid = lambda w, h, tile_size=96: orig(w, h, tile_size)
    processor.set_tile_mode(True)
    iters = processor.core.model.iters
    processor.core.model.iters = 4
    try:
        full = processor.compute_optical_flow_tiled(frames, 2)
        assert full.shape == (192, 352, 2)
        processor.set_tile_mode(False)
        for ti in orig(352, 192, 96)[4]:
            x, y, w, h = ti['x'], ti['y'], ti['width'], ti['height']
            tile = processor.compute_optical_flow([f[y:y + h, x:x + w] for f in frames], 2)
            # not bit-equal: cuDNN/cuBLAS in the torch encoder are not run-to-run identical across
            # differently-placed inputs (fmaps differ by ~4e-6), the libvfml kernels are deterministic
            assert np.abs(full[y:y + h, x:x + w] - tile).max() < 1e-3
    finally:
        processor.calculate_tile_grid = orig
        processor.set_tile_mode(False)
        processor.core.model.iters = iters


def test_post_pass_wrappers(golden):
    from video_flow_ml_b200.effects import TAAProcessor
    from video_flow_ml_b200.encoding import encode_flow, FlowEncoderFactory
    g = golden("post.npz")
    ff = cases.flow_field_case()
    assert np.array_equal(encode_flow(ff.copy(), ff.shape[1], ff.shape[0], "gamedev"), g["gamedev"])
    enc = FlowEncoderFactory.create_encoder("motion-vectors-rg8", clamp_range=16.0)
    assert np.array_equal(enc.encode(ff, 0, 0), g["rg8_c16"])
    with pytest.raises(ValueError):
        FlowEncoderFactory.create_encoder("nope")
    frames, flows = cases.taa_case()
    for name, kw in (("taa_bilateral", dict(use_flow=True, use_bilateral=True)),
                     ("taa_bilinear", dict(use_flow=True, use_bilateral=False)),
                     ("taa_simple", dict(use_flow=False, use_bilateral=False))):
        t = TAAProcessor(alpha=0.1)
        prev = None
        for i, f in enumerate(frames):
            o = t.apply_taa(f, prev, alpha=0.1, sequence_id="s", **kw)
            want = g[f"{name}_{i}"]
            assert o.dtype == want.dtype, (name, i, o.dtype, want.dtype)
            assert np.abs(o.astype(np.float64) - want).max() <= 1e-3, (name, i)
            prev = flows[i].copy()


def test_mof_fp16_tier_matches_reference_golden(golden):
    """Tensor-core tier (fp16 operands, fp32 accumulation, like the reference's own fp16 autocast on CUDA):
    max EPE delta <= 5e-2 px vs the fp32 output of the unmodified reference (north-star reduced-precision bound)."""
    from video_flow_ml_b200.processing import VideoFlowProcessor
    g = golden("mof_e2e_g10.npz")
    p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp16")
    p.core.load_state_dict(omof.mof_synthetic_weights(0, gain=1.0))
    frames = cases.mof_case()
    for idx in (2, 5):
        flow = p.compute_optical_flow(list(frames), idx)
        assert epe(flow, g[f"flow_{idx}"]) <= 5e-2, (idx, epe(flow, g[f"flow_{idx}"]))


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("fp16", 5e-2)])
def test_bof_matches_reference_golden(golden, precision, tol):
    """BOF architecture (3 frames, 12 iterations) against the unmodified reference BOFNet; the wrapper returns
    the 2->1 flow (index shape[1] // 2 of the model output, processing/videoflow_core.py:192-194)."""
    from oracle import bof as obof
    from video_flow_ml_b200.processing import VideoFlowProcessor
    g = golden("bof_e2e.npz")
    p = VideoFlowProcessor("cuda", sequence_length=3, architecture="bof", dataset="things", precision=precision)
    p.core.load_state_dict(obof.bof_synthetic_weights(0))
    frames = cases.mof_case()
    for idx in (2, 4):
        flow = p.compute_optical_flow(list(frames), idx)
        assert epe(flow, g[f"flows_{idx}"][1].transpose(1, 2, 0)) <= tol, (idx, epe(flow, g[f"flows_{idx}"][1].transpose(1, 2, 0)))
    # all flows of the window: [2->3, 2->1]
    win = p.prepare_frame_sequence(list(frames), 2)
    x = torch.nn.functional.pad(win, [2, 2, 2, 2, 0, 0], mode="replicate")
    allf = p.core.model.forward(x)[0, :, :, 2:-2, 2:-2].cpu().numpy()
    assert np.sqrt(((allf - g["flows_2"]) ** 2).sum(1)).max() <= tol


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("fp16", 5e-2)])
def test_memflow_matches_reference_golden(golden, precision, tol):
    """MemFlow (15 iterations) against the unmodified reference MemFlowNet + InferenceCore: a 3-step sequence through
    the stateful InferenceCore mirror (memory empty / one stored frame / replaced), then the application's call
    (MemFlowCore.compute_flow_from_tensor: fresh memory, last two frames, end=True, unpad)."""
    from oracle import memflow as omem
    from video_flow_ml_b200.processing import InferenceCore, MemFlowCore, MemFlowProcessor
    g = golden("memflow_e2e.npz")
    core = MemFlowCore("cuda", precision=precision)
    with pytest.raises(RuntimeError):
        core.compute_flow_from_tensor(torch.zeros(1, 2, 3, 64, 64))
    with pytest.raises(FileNotFoundError):
        core.load_model()
    core.load_state_dict({"module." + k: v for k, v in omem.memflow_synthetic_weights(0).items()})
    frames = cases.mof_case()
    x = torch.stack([torch.from_numpy(f).permute(2, 0, 1) for f in frames[:4]])[None]
    x = omof.pad_frames(2 * (x / 255.0) - 1.0).contiguous().cuda()
    ic = InferenceCore(core.model)
    for rep in range(2):                    # second pass: cleared memory, replayed graphs
        ic.clear_memory()
        for t in range(3):
            low, up = ic.step(x[:, t:t + 2])
            e_low = epe(low[0].permute(1, 2, 0).cpu().numpy(), g[f"seq_low_{t}"][0].transpose(1, 2, 0))
            e_up = epe(up[0].permute(1, 2, 0).cpu().numpy(), g[f"seq_up_{t}"][0].transpose(1, 2, 0))
            assert e_up <= tol and e_low <= tol / 8 + 1e-3, (rep, t, e_low, e_up)
            assert core.model.n_stored == 1
    win = omof.frames_to_window(frames, 2, 3)
    flow = core.compute_flow_from_tensor(win)
    assert flow.shape == (2, 124, 188) and flow.device.type == "cpu"
    assert epe(flow.permute(1, 2, 0).numpy(), g["core_flow"].transpose(1, 2, 0)) <= tol
    for bad, exc in ((torch.zeros(1, 1, 3, 64, 64), ValueError), (torch.zeros(2, 2, 3, 64, 64), ValueError),
                     (torch.zeros(1, 2, 3, 32, 64), ValueError), (np.zeros((1, 2, 3, 64, 64)), TypeError)):
        with pytest.raises(exc):
            core.compute_flow_from_tensor(bad)
    # processor plumbing: uint8 frames, numpy [H, W, 2] out, same value as the core on the same window
    p = MemFlowProcessor("cuda", precision=precision)
    p.core_engine = core
    u8 = [np.clip(f, 0, 255).astype(np.uint8) for f in frames]
    out = p.compute_optical_flow(u8, 3)
    assert out.shape == (124, 188, 2) and out.dtype == np.float32
    want = core.compute_flow_from_tensor(p.prepare_frame_sequence(u8, 3)).permute(1, 2, 0).numpy()
    assert np.array_equal(out, want)
    assert p.calculate_tile_grid(188, 124)[2:4] == (1, 1)


def test_device_resident_post_pass(processor_g10):
    """flow -> gamedev encode -> TAA without the D2H -> numpy hop of flow_processor.py:976-1112 gives
    bit-identical results to the host-array path (the kernels are the same; only the copies disappear)."""
    from video_flow_ml_b200.effects import TAAProcessor
    from video_flow_ml_b200.encoding import encode_flow
    frames = [np.clip(f, 0, 255).astype(np.uint8) for f in cases.mof_case()]
    H, W = frames[0].shape[:2]
    host_taa, dev_taa = TAAProcessor(alpha=0.1), TAAProcessor(alpha=0.1)
    for idx in (2, 3):
        flow_np = processor_g10.compute_optical_flow(frames, idx)
        flow_dev = processor_g10.compute_optical_flow(frames, idx, return_device=True)
        assert flow_dev.is_cuda and flow_dev.shape == (H, W, 2)
        assert np.array_equal(flow_dev.cpu().numpy(), flow_np)
        enc_np = encode_flow(flow_np, W, H, 'gamedev')
        enc_dev = encode_flow(flow_dev, W, H, 'gamedev')
        assert enc_dev.is_cuda and enc_dev.dtype == torch.uint8 and np.array_equal(enc_dev.cpu().numpy(), enc_np)
        out_np = host_taa.apply_taa(enc_np, flow_np)
        out_dev = dev_taa.apply_taa(enc_dev, flow_dev, return_device=True)
        assert out_dev.is_cuda and np.array_equal(out_dev.cpu().numpy().astype(out_np.dtype), out_np)


@pytest.mark.parametrize("T,H,W,precision,tol", [(7, 124, 188, "fp32", 1e-2), (9, 124, 252, "fp32", 1e-2),
                                                 (7, 124, 188, "fp16", 5e-2)])
def test_mof_long_sequences_match_oracle(T, H, W, precision, tol):
    """Sequence lengths 7 and 9 (BASELINE.json configs[4] / configs[2]) through `set_sequence_length`
    (videoflow_processor.py:191-195): window choice, T-2 centre frames, returned flow index T-2, against the pinned
    oracle on the same frames and weights (well-conditioned weight set).  Tolerances: the north star's."""
    from video_flow_ml_b200.processing import VideoFlowProcessor
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    p = VideoFlowProcessor("cuda", sequence_length=5, precision=precision)
    p.set_sequence_length(T)
    p.core.load_state_dict(sd, iters=8)
    frames = cases.mof_case(seed=77, n_frames=T + 1, H=H, W=W)
    idx = T // 2
    flow = p.compute_optical_flow(list(frames), idx)
    assert flow.shape == (H, W, 2) and flow.dtype == np.float32
    win = omof.frames_to_window(frames, idx, T)
    want = omof.core_compute_flow(win, sd, iters=8).permute(1, 2, 0).numpy()     # flow index T-2
    assert epe(flow, want) <= tol, epe(flow, want)


# ---------------------------------------------------------------------------- round 2: large motion, BASELINE sizes, tiles
def _proc(precision, sd, T=5, iters=32, arch="mof", dataset="sintel"):
    from video_flow_ml_b200.processing import VideoFlowProcessor
    p = VideoFlowProcessor("cuda", sequence_length=T, precision=precision, architecture=arch, dataset=dataset)
    p.core.load_state_dict(sd, iters=iters)
    return p


@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_mof_large_motion_fixture(golden, precision):
    """Large-motion fixture (tests/golden/make_golden_r2.py `drift`): gain-1.0 weights + a drift in the flow-head bias,
    the low-res coordinates travel ~20 texels (|flow| up to 162 px) and leave the image, so the 64 recurrent lookups of
    a window run far from the identity grid.  The fixture's own conditioning is stored with it: a 1-ulp perturbation
    of the encoder outputs moves the REFERENCE's output by 1.0e-3 px, and the reference's own shipped mixed-precision
    path (fp16 autocast, fp32 volume + lookup; emulated on the pinned oracle) deviates by 1.64 px from its fp32 output:
    a +-20-texel flow value has an fp16 ulp of 1.6e-2 texels, so NO fp16 path can hold 5e-2 px here.
      fp32 tier: <= 1e-2 px (north-star fp32 tolerance; measured 1.6e-3).
      fp16 tier: no further from the reference's fp32 output than the reference's OWN mixed-precision path is
                 (floor_fp16 = 1.64 px; measured 1.32 px max, and profiles/r2_fp16_ablation.txt shows that flipping any
                 single reduced-precision element back to fp32 -- lookup arithmetic, volume storage, encoders, MUFU GELU --
                 only moves the result inside that noise: 0.89 .. 1.34 px).  The 5e-2 px bound is asserted where an fp16
                 path can meet it: mof_e2e_g10, the C1 / C3 / BOF-tile / tiled fixtures."""
    g = golden("mof_e2e_drift.npz")
    sd = omof.mof_synthetic_weights(0, gain=1.0, drift=omof.DRIFT)
    p = _proc(precision, sd)
    frames = cases.mof_case()
    tol = 1e-2 if precision == "fp32" else float(g["floor_fp16"])
    for idx in (2, 5):
        flow = p.compute_optical_flow(list(frames), idx)
        assert np.abs(g[f"flow_{idx}"]).max() > 100.0
        d = np.sqrt(((flow - g[f"flow_{idx}"]) ** 2).sum(-1))
        e = d.max()
        print(f"large-motion fixture, {precision}, frame {idx}: EPE delta max {e:.3e} / mean {d.mean():.3e} / p99 {np.percentile(d, 99):.3e} px "
              f"(bound {tol:.3e}), |flow| max {np.abs(g[f'flow_{idx}']).max():.0f} px")
        assert e <= tol, (idx, e)


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("fp16", 5e-2)])
def test_c1_size_matches_reference_golden(golden, precision, tol):
    """BASELINE config C1 size (436x1024, T=5 -> 55x128 features, P = 7040) against the unmodified reference wrapper
    (mof_c1.npz, 4 iterations, drift weights: the coordinates move 2.4 texels): the paths that only trigger at size
    (weight-stationary GEMM mode, wave quantisation, split-K aggregate, 32-row depthwise tiles, PDL) are parity-checked."""
    g = golden("mof_c1.npz")
    sd = omof.mof_synthetic_weights(0, gain=1.0, drift=omof.DRIFT)
    p = _proc(precision, sd, iters=4)
    frames = [f.astype(np.float32) for f in cases.synthetic_clip(8, 436, 1024, seed=1234)]
    flow = p.compute_optical_flow(frames, 3)
    assert flow.shape == (436, 1024, 2)
    e = epe(flow[::3, ::3], g["flow_3"])
    print(f"C1 size, {precision}: max EPE delta {e:.3e} px, |flow| max {np.abs(g['flow_3']).max():.1f}")
    assert e <= tol, e


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("fp16", 5e-2)])
def test_c3_size_matches_reference_golden(golden, precision, tol):
    """BASELINE config C3 size: 376x1242 (-> 376x1248, 47x156 features), T=9 (7 centre frames), 4 iterations."""
    g = golden("mof_c3.npz")
    sd = omof.mof_synthetic_weights(0, gain=1.0, drift=omof.DRIFT)
    p = _proc(precision, sd, T=9, iters=4)
    frames = [f.astype(np.float32) for f in cases.synthetic_clip(10, 376, 1242, seed=4321)]
    flow = p.compute_optical_flow(frames, 4)
    e = epe(flow[::3, ::3], g["flow_4"])
    print(f"C3 size, {precision}: max EPE delta {e:.3e} px")
    assert e <= tol, e


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("fp16", 5e-2)])
def test_bof_1080p_tile_matches_reference_golden(golden, precision, tol):
    """BASELINE config C2: the 1280x1080 tile of a 1080p frame (P = 21600) through BOF, 4 iterations, against the
    unmodified BOFNet (bof_tile.npz)."""
    from oracle import bof as obof
    g = golden("bof_tile.npz")
    p = _proc(precision, obof.bof_synthetic_weights(0), T=3, iters=4, arch="bof", dataset="things")
    frames = [f.astype(np.float32)[:, :1280] for f in cases.synthetic_clip(3, 1080, 1920, seed=99)]
    flow = p.compute_optical_flow(frames, 1)                       # the 2->1 flow = index 1 of the model output
    e = epe(flow[::5, ::5], g["flows_1"][1].transpose(1, 2, 0))
    print(f"BOF 1280x1080 tile, {precision}: max EPE delta {e:.3e} px")
    assert e <= tol, e
    del p
    torch.cuda.empty_cache()


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("fp16", 5e-2)])
def test_tiled_matches_reference_golden(golden, precision, tol):
    """compute_optical_flow_tiled against the REFERENCE's compute_optical_flow_tiled (mof_tiled.npz; tile size 160 through
    calculate_tile_grid's parameter on a 288x448 clip: 2x3 ragged tiles of 160/128 px), uint8 and float frames, a
    clipped window (frame 0), with and without the per-frame feature cache."""
    g = golden("mof_tiled.npz")
    p = _proc(precision, omof.mof_synthetic_weights(0, gain=1.0), iters=4)
    orig = p.calculate_tile_grid
    p.calculate_tile_grid = lambda w, h, tile_size=160: orig(w, h, tile_size)
    p.set_tile_mode(True)
    frames = cases.synthetic_clip(5, 288, 448, seed=5)
    ffl = [f.astype(np.float32) for f in frames]
    for cache in (True, False):
        p.set_feature_cache(cache)
        full = p.compute_optical_flow_tiled(list(ffl), 2)
        assert full.shape == (288, 448, 2)
        assert epe(full, g["tiled_f32_2"]) <= tol, (cache, epe(full, g["tiled_f32_2"]))
        assert epe(p.compute_optical_flow_tiled(list(ffl), 0)[::3, ::3], g["tiled_f32_0"]) <= tol
        assert epe(p.compute_optical_flow_tiled(list(frames), 2)[::3, ::3], g["tiled_u8_2"]) <= tol
    assert len(p.core.model._ws) >= 2             # ragged grid: two window shapes stay resident (no re-capture per tile)


def test_feature_cache_equals_uncached_and_saves_encoder_passes():
    """Sliding windows through the per-frame feature cache: same flows as the self-contained tensor path (encoders are
    per-image functions; cuDNN picks batch-dependent kernels, hence 1e-3 px rather than bit equality), one new fnet and
    one new cnet pass per steady-state window."""
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    p = _proc("fp32", sd, iters=4)
    p.core.prefetch = False          # (the prefetch of the next window's frame would add its encoder passes to the counts)
    frames = [np.clip(f, 0, 255).astype(np.uint8) for f in cases.mof_case(seed=31, n_frames=8)]
    got = [p.compute_optical_flow(frames, i) for i in range(2, 6)]
    st = dict(p.core.cache_stats)
    assert st["fnet_runs"] == 5 + 3 and st["cnet_runs"] == 3 + 3, st
    p.set_feature_cache(False)
    for i, g_ in zip(range(2, 6), got):
        want = p.compute_optical_flow(frames, i)
        assert np.abs(want - g_).max() <= 1e-3, (i, np.abs(want - g_).max())
    assert p.core.cache_stats == st
    # a frame rewritten in place is noticed: its features are recomputed (uint8 frames take the reference's double
    # normalisation, SURVEY Q1, so the flow itself barely depends on the content: the encoder pass count is the evidence)
    p.set_feature_cache(True)
    p.compute_optical_flow(frames, 3)
    before = dict(p.core.cache_stats)
    p.compute_optical_flow(frames, 3)
    assert p.core.cache_stats["fnet_runs"] == before["fnet_runs"]            # unchanged frames: all hits
    frames[3][:] = frames[3][::-1].copy()
    p.compute_optical_flow(frames, 3)
    assert p.core.cache_stats["fnet_runs"] == before["fnet_runs"] + 1 and p.core.cache_stats["cnet_runs"] == before["cnet_runs"] + 1


def test_feature_cache_arena_recycles_slots_without_allocating():
    """The cached per-frame features live in a fixed arena (one allocation per shape): a cache of two windows (10 slots)
    sliding over 24 frames re-uses evicted frames' slots -- bit-identical flows to the default 96-slot cache, which never
    evicts here -- and once the arena exists a window takes no memory from the driver (cudaMalloc stalls the launching
    thread on the B200 VMs)."""
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    frames = [np.ascontiguousarray(f) for f in cases.sliding_clip(0, 24, 64, 96, seed=5)]
    p = _proc("fp16", sd, iters=3)
    want = [p.compute_optical_flow(frames, i) for i in range(len(frames))]
    arena = next(iter(p.core._arenas.values()))
    assert arena.cap == 96 and len(arena.free) == 96 - 24
    p.core.clear_feature_cache()
    p.core.feature_cache_frames = 10
    reserved = []
    for i in range(len(frames)):
        assert np.array_equal(p.compute_optical_flow(frames, i), want[i]), i
        reserved.append(torch.cuda.memory_reserved())
    arena = next(iter(p.core._arenas.values()))
    assert arena.cap == 10 and len(p.core._feat) == 10 and len(arena.free) == 0
    assert sorted(e["slot"] for e in p.core._feat.values()) == list(range(10))
    assert reserved[-1] == reserved[8], (reserved[8], reserved[-1])        # steady state: nothing new from cudaMalloc
    # a cache smaller than two windows is not pooled (a slot could be recycled under the window that still reads it)
    p.core.clear_feature_cache()
    p.core.feature_cache_frames = 4
    assert np.array_equal(p.compute_optical_flow(frames, 12), want[12]) and not p.core._arenas


def test_mof_fp32_split_tier_matches_reference_golden(golden):
    """Opt-in tensor-core route of the fp32 tier (ops.set_fp32_split: fp32 products as three fp16 tensor-core products):
    within 1e-3 px of the REFERENCE on the gain-1.0 fixture (measured 1.2e-4; the FFMA tier: 3.6e-5)."""
    from video_flow_ml_b200 import ops
    g = golden("mof_e2e_g10.npz")
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    ops.set_fp32_split(True)
    try:
        p = _proc("fp32", sd)
        frames = cases.mof_case()
        for idx in (2, 5):
            flow = p.compute_optical_flow(list(frames), idx)
            assert epe(flow, g[f"flow_{idx}"]) <= 1e-3, (idx, epe(flow, g[f"flow_{idx}"]))
    finally:
        ops.set_fp32_split(False)


def test_pyramid_reuse_between_sliding_windows_is_bit_exact():
    """A window that follows its predecessor by one frame re-uses the correlation pyramids of the shared centre frames
    and their attention matrices, in place (ring order: one new pair instead of 2 (T-2) pairs, one CUDA graph per rotation):
    bit-identical flows with and without
    (same call sequence, so the per-frame encoder cache behaves identically in both runs), also across jumps in the clip
    (re-use must not trigger) and after an in-place edit of a shared frame (its pairs are rebuilt)."""
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    orig = [f.astype(np.float32) for f in cases.mof_case(seed=77, n_frames=11)]
    order = [2, 3, 4, 5, 8, 9, 3, 4]                 # slides, a jump, a slide, a jump back, a slide

    def run(cache, prefetch=True, ring=True):
        frames = [f.copy() for f in orig]
        p = _proc("fp16", sd, iters=3)
        eng = p.core.model
        assert eng.tiled
        p.core.prefetch = prefetch
        eng.pyramid_cache = cache
        eng.pyramid_ring = ring
        got = [p.compute_optical_flow(frames, i) for i in order]
        slides = [eng.pyramid_slides]                               # windows 3, 4, 5 | 9 | 4
        frames[6][:] = frames[6][:, ::-1].copy()                    # frame 6 is in two pairs the window at 5 would have re-used
        got.append(p.compute_optical_flow(frames, 5))
        slides.append(eng.pyramid_slides)
        got.append(p.compute_optical_flow(frames, 6))
        slides.append(eng.pyramid_slides)
        return got, slides

    got, slides = run(True)
    want, none = run(False)
    assert slides == [5, 5, 6] and none == [0, 0, 0], (slides, none)
    for k, (a, b) in enumerate(zip(want, got)):
        assert np.array_equal(a, b), (k, float(np.abs(a - b).max()))
    # the encoder prefetch of the next window's new frame (side stream, beside the iterations) changes nothing either
    plain, _ = run(False, prefetch=False)
    for k, (a, b) in enumerate(zip(plain, got)):
        assert np.array_equal(a, b), ("prefetch", k, float(np.abs(a - b).max()))
    # ring order off: pyramids moved between slots by device copies, attention recomputed every window
    moved, slides = run(True, ring=False)
    assert slides == [5, 5, 6]
    for k, (a, b) in enumerate(zip(moved, got)):
        assert np.array_equal(a, b), ("ring", k, float(np.abs(a - b).max()))


def test_ring_order_long_windows_every_rotation():
    """T = 7 (five centre frames): eleven consecutive windows take the ring through every rotation twice; each flow equals the
    one a fresh pass computes for that window alone (same per-frame features through the cache: bit-identical)."""
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    frames = [f.astype(np.float32) for f in cases.mof_case(seed=78, n_frames=18)]
    p = _proc("fp16", sd, T=7, iters=2)
    got = [p.compute_optical_flow(frames, i) for i in range(3, 15)]
    assert p.core.model.pyramid_slides == 11
    ws = next(iter(p.core.model._ws.values()))
    assert ws["rot"] == 11 % 5 and len(ws["graphs"]) == 6          # one cold graph + one per rotation
    p.core.model.pyramid_cache = False
    for i, g_ in zip(range(3, 15), got):
        assert np.array_equal(p.compute_optical_flow(frames, i), g_), i


def test_device_resident_post_pass_matches_reference_golden(golden):
    """SURVEY 8(f) n2 against the REFERENCE's outputs (post.npz), not against this repo's host path: the device-resident
    encode and TAA (CUDA tensors in, CUDA tensors out, no host hop) reproduce the reference encoders bit for bit and
    its TAAProcessor within 1e-3."""
    from video_flow_ml_b200.effects import TAAProcessor
    from video_flow_ml_b200.encoding import encode_flow
    g = golden("post.npz")
    ff = torch.from_numpy(cases.flow_field_case()).cuda()
    enc = encode_flow(ff, ff.shape[1], ff.shape[0], "gamedev")
    assert enc.is_cuda and np.array_equal(enc.cpu().numpy(), g["gamedev"])
    frames, flows = cases.taa_case()
    t = TAAProcessor(alpha=0.1)
    prev = None
    for i, f in enumerate(frames):
        o = t.apply_taa(torch.from_numpy(f).cuda(), prev, alpha=0.1, sequence_id="s", return_device=True)
        assert o.is_cuda and np.abs(o.cpu().numpy() - g[f"taa_bilateral_{i}"]).max() <= 1e-3, i
        prev = torch.from_numpy(flows[i]).cuda()


def test_real_4k_tile_runs_and_fits():
    """One real 1280x1280 tile at T=7 (BASELINE config C5: P = 25600, 5 centre frames, 10 frame pairs): the fp16 tier's
    workspace fits, the size-dependent kernel choices hold at P = 25600, and its flow agrees with the exact fp32 tier
    (2 iterations; the CPU oracle at this size needs ~50 GB and minutes, so the fp32 tier -- itself pinned to the
    reference at every other size -- stands in)."""
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    frames = [f.astype(np.float32) for f in cases.synthetic_clip(7, 1280, 1280, seed=77)]
    flows = {}
    for prec in ("fp16", "fp32"):
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
        p = _proc(prec, sd, T=7, iters=2)
        flows[prec] = p.compute_optical_flow(frames, 3)
        print(f"1280x1280 T=7 tile, {prec}: peak device memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
        assert torch.cuda.max_memory_allocated() < 120 * 2**30
        del p
    assert np.isfinite(flows["fp16"]).all() and flows["fp16"].shape == (1280, 1280, 2)
    e = epe(flows["fp16"], flows["fp32"])
    print(f"1280x1280 T=7 tile: fp16 vs fp32 tier max EPE delta {e:.3e} px")
    assert e <= 5e-2, e


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_non_default_device():
    """A processor built for 'cuda:1' works without torch.cuda.set_device(1) (the reference wrappers accept any device)."""
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    from video_flow_ml_b200.processing import VideoFlowProcessor
    from video_flow_ml_b200.encoding import encode_flow
    frames = cases.mof_case()
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        p = VideoFlowProcessor(dev, sequence_length=5, precision="fp16")
        p.core.load_state_dict(sd, iters=4)
        fl = p.compute_optical_flow(list(frames), 2, return_device=True)
        assert str(fl.device) == dev
        enc = encode_flow(fl.contiguous(), 188, 124, "gamedev")
        assert str(enc.device) == dev
        outs.append(fl.cpu().numpy())
    assert torch.cuda.current_device() == 0
    assert np.abs(outs[0] - outs[1]).max() <= 1e-3
