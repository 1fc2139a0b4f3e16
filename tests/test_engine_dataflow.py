"""Host-side orchestration of the MOF engine (buffer carving, column ranges, call order, weight
prepacking) against the oracle, with the libvfml ops replaced by the CPU test double."""
import numpy as np
import pytest
import torch

import cases
import fake_ops
from oracle import mof as omof
from oracle import twins as otwins
from video_flow_ml_b200.mof import MOFEngine
from video_flow_ml_b200.twins import TwinsStage01

torch.set_grad_enabled(False)


def test_twins_matches_oracle(mof_weights):
    x = torch.from_numpy(cases.twins_case())
    for pre in ("fnet.svt", "cnet.svt"):
        got = TwinsStage01(mof_weights, pre, "cpu")(x)
        want = otwins.twins_svt_forward(x, mof_weights, pre)
        assert (got - want).abs().max() < 1e-4


def test_engine_dataflow_matches_oracle(mof_weights):
    frames = cases.mof_case()
    win = omof.pad_frames(omof.frames_to_window(frames, 2, 5)).contiguous()
    eng = MOFEngine(mof_weights, "cpu", torch.float32, iters=3, ops=fake_ops)
    fmaps, net, inp = eng.encode(win)
    of, on, oi = omof.mof_encode(win, mof_weights)
    assert (fmaps - of).abs().max() < 1e-4 and (net - on).abs().max() < 1e-4 and (inp - oi).abs().max() < 1e-4
    want = omof.mof_refine(of, on, oi, mof_weights, iters=3)
    got = eng.refine(of, on, oi)
    assert got.shape == want.shape
    assert (got - want).abs().max() < 2e-3, (got - want).abs().max()
    sel = eng.refine(of, on, oi, select=3, crop=(2, 1, 120, 180))
    assert (sel - want[0, 3, :, 2:122, 1:181]).abs().max() < 2e-3


def test_engine_tiled_lookup_channel_permutation(mof_weights):
    """The tiled lookup emits the 324 correlation channels in a permuted order; the engine folds the
    permutation into the convc1 weights.  Result must equal the unpermuted path."""
    frames = cases.mof_case()
    win = omof.pad_frames(omof.frames_to_window(frames, 2, 5)).contiguous()
    of, on, oi = omof.mof_encode(win, mof_weights)
    ref = MOFEngine(mof_weights, "cpu", torch.float32, iters=2, ops=fake_ops, tiled=False).refine(of, on, oi)
    got = MOFEngine(mof_weights, "cpu", torch.float32, iters=2, ops=fake_ops, tiled=True).refine(of, on, oi)
    assert (got - ref).abs().max() < 1e-3, (got - ref).abs().max()      # fp32 summation order differs
    perm = fake_ops.corr_lookup_perm()
    assert sorted(perm.tolist()) == list(range(324))
    assert perm[0] == 0 and perm[7] == 7 and perm[8] == 9 and perm[288] == 8 and perm[323] == 3 * 81 + 8 * 9 + 8


def test_bof_engine_dataflow_matches_oracle():
    """BOF (3-frame) variant of the engine: no hidden state, 256-channel conv block, [2->1 | 2->3] mask order."""
    from oracle import bof as obof
    sd = obof.bof_synthetic_weights(0)
    frames = cases.mof_case()
    win = omof.pad_frames(omof.frames_to_window(frames, 2, 3)).contiguous()
    of, on, oi = obof.bof_encode(win, sd)
    want = obof.bof_refine(of, on, oi, sd, iters=3)
    eng = MOFEngine(sd, "cpu", torch.float32, iters=3, ops=fake_ops, arch="bof")
    got = eng.refine(of, on, oi)
    assert got.shape == want.shape == (1, 2, 2, 128, 192)
    assert (got - want).abs().max() < 2e-3, (got - want).abs().max()
    sel = eng.refine(of, on, oi, select=1)          # what VideoFlowCore returns: the 2->1 flow
    assert (sel - want[0, 1]).abs().max() < 2e-3


def test_memflow_engine_dataflow_matches_oracle():
    """MemFlow engine (memory slots, once-per-frame affinity, 4-column glue trick) on CPU doubles vs the oracle's
    InferenceCore: three steps so the memory is empty, then engaged, then replaced; plus end=True and flow_init."""
    from oracle import memflow as omem
    from video_flow_ml_b200.memflow import MemFlowEngine
    sd = omem.memflow_synthetic_weights(0)
    frames = cases.mof_case()
    x = torch.stack([torch.from_numpy(f).permute(2, 0, 1) for f in frames[:4]])[None]
    x = omof.pad_frames(2 * (x / 255.0) - 1.0).contiguous()
    want_core = omem.InferenceCore(sd, iters=3)
    for tiled in (False, True):
        eng = MemFlowEngine(sd, "cpu", torch.float32, config={"val_decoder_depth": 3}, ops=fake_ops, tiled=tiled)
        want_core.clear_memory()
        for t in range(3):
            wl, wu = want_core.step(x[:, t:t + 2])
            gl, gu = eng.step(x[:, t:t + 2])
            assert gl.shape == wl.shape == (1, 2, 16, 24) and gu.shape == wu.shape == (1, 2, 128, 192)
            assert (gl - wl).abs().max() < 2e-4 and (gu - wu).abs().max() < 2e-3, (t, (gu - wu).abs().max())
            assert eng.n_stored == 1
    # end=True never memorises; flow_init shifts the starting coordinates (MemFlow.py:95-96)
    eng.clear_memory()
    want_core.clear_memory()
    fi = 0.5 * torch.ones(1, 2, 16, 24)
    wl, wu = want_core.step(x[:, 1:3], end=True, flow_init=fi)
    gl, gu = eng.step(x[:, 1:3], end=True, flow_init=fi)
    assert eng.n_stored == 0 and want_core.mem_k is None
    assert (gu - wu).abs().max() < 2e-3
    with pytest.raises(ValueError):
        eng.step(x[:, 0:3])


@pytest.mark.parametrize("max_f,min_f,every", [(3, 2, 1), (2, 1, 2), (4, 1, 1)])
def test_memflow_memory_policy_variants(max_f, min_f, every):
    """Working-memory bookkeeping for other configs than the shipped 2 / 1 / 1 (memory_manager_skflow.py:99-124,
    inference_core_skflow.py:24): frames stored, compression to the newest `min` frames, mem_every."""
    from oracle import memflow as omem
    from video_flow_ml_b200.memflow import MemFlowEngine
    sd = omem.memflow_synthetic_weights(0)
    frames = cases.mof_case()
    x = torch.stack([torch.from_numpy(f).permute(2, 0, 1) for f in frames[:6]])[None]
    x = omof.pad_frames(2 * (x / 255.0) - 1.0).contiguous()
    want = omem.InferenceCore(sd, iters=2, mem_every=every, max_frames=max_f, min_frames=min_f)
    eng = MemFlowEngine(sd, "cpu", torch.float32, ops=fake_ops, tiled=False,
                        config={"val_decoder_depth": 2, "mem_every": every, "max_mid_term_frames": max_f, "min_mid_term_frames": min_f})
    P = 16 * 24
    for t in range(5):
        wl, wu = want.step(x[:, t:t + 2])
        gl, gu = eng.step(x[:, t:t + 2])
        assert (gu - wu).abs().max() < 2e-3, (t, (gu - wu).abs().max())
        stored = 0 if want.mem_k is None else want.mem_k.shape[-1] // P
        assert eng.n_stored == stored, (t, eng.n_stored, stored)


@pytest.mark.parametrize("T,H,W", [(7, 124, 188), (9, 124, 252)])
def test_engine_dataflow_long_sequences(mof_weights, T, H, W):
    """Sequence lengths 7 and 9 (BASELINE.json configs[4] and configs[2]; `set_sequence_length`,
    videoflow_processor.py:191-195) on frame sizes that need padding (a pyramid level of height 1 makes the reference itself return NaN: bilinear_sampler divides by H-1, so 96-row inputs are avoided):
    T-2 centre frames, 2(T-2) flows, the temporal hidden-state warp reaches across every neighbour pair."""
    frames = cases.mof_case(seed=77, n_frames=T + 1, H=H, W=W)
    win = omof.pad_frames(omof.frames_to_window(frames, T // 2, T)).contiguous()
    assert win.shape[1] == T and win.shape[-2] % 8 == 0 and win.shape[-1] % 8 == 0
    of, on, oi = omof.mof_encode(win, mof_weights)
    want = omof.mof_refine(of, on, oi, mof_weights, iters=2)
    eng = MOFEngine(mof_weights, "cpu", torch.float32, iters=2, ops=fake_ops)
    got = eng.refine(of, on, oi)
    assert got.shape == want.shape == (1, 2 * (T - 2), 2, win.shape[-2], win.shape[-1])
    assert (got - want).abs().max() < 2e-3, (got - want).abs().max()
    mid = (T - 2) // 2                               # the flow VideoFlowCore returns: centre frame -> next
    sel = eng.refine(of, on, oi, select=mid)
    assert (sel - want[0, mid]).abs().max() < 2e-3


def test_pyramid_pair_keys_identify_shared_pyramids():
    """MOFEngine._pair_keys: (centre, partner) identity of every pyramid slot; a window shifted by one frame shares all
    but the last forward and backward pair with its predecessor, a jump or an edited frame does not."""
    from video_flow_ml_b200.mof import MOFEngine
    Fc = 3
    a = MOFEngine._pair_keys(list("abcde"), Fc)
    b = MOFEngine._pair_keys(list("bcdef"), Fc)
    assert a == (("b", "c"), ("c", "d"), ("d", "e"), ("b", "a"), ("c", "b"), ("d", "c"))
    shared = lambda new, prev: all(new[j] == prev[j + 1] and new[Fc + j] == prev[Fc + j + 1] for j in range(Fc - 1))
    assert shared(b, a)
    assert not shared(MOFEngine._pair_keys(list("cdefg"), Fc), a)            # a jump of two frames
    assert not shared(MOFEngine._pair_keys(["b", "c", "D", "e", "f"], Fc), a)   # frame d was edited (new key)
    assert MOFEngine._pair_keys(["a", None, "c", "d", "e"], Fc) is None and MOFEngine._pair_keys(None, Fc) is None


@pytest.mark.parametrize("ring", [True, False])
def test_engine_sliding_windows_keep_per_frame_state_in_ring_order(mof_weights, ring):
    """Consecutive windows given with frame identities (refine(frame_keys=...)): the engine keeps the pyramids -- and, in ring
    mode, the attention matrices -- of the centre frames two windows share, computes only the new frame's, and in ring mode
    rotates the slot order instead of moving data (motion_glue's temporal neighbours, the context rows and the selected output
    follow the rotation).  Every window must equal a fresh engine's self-contained result; seven slides cover every rotation
    of Fc = 3 twice; a jump and a repeated window fall back to the full build."""
    torch.manual_seed(0)
    frames = cases.mof_case(seed=5, n_frames=12)
    eng = MOFEngine(mof_weights, "cpu", torch.float32, iters=2, ops=fake_ops, tiled=True)
    eng.pyramid_ring = ring
    order = [2, 3, 4, 5, 6, 7, 8, 9, 4, 5, 5, 6]
    rots = []
    for n, i in enumerate(order):
        win = omof.pad_frames(omof.frames_to_window(frames, i, 5)).contiguous()
        of, on, oi = omof.mof_encode(win, mof_weights)
        keys = [("f", k) for k in range(i - 2, i + 3)]
        fresh = MOFEngine(mof_weights, "cpu", torch.float32, iters=2, ops=fake_ops, tiled=True).refine(of, on, oi)
        if n % 2:                                                 # all flows, in window order whatever the rotation
            got = eng.refine(of, on, oi, frame_keys=keys)
            assert (got - fresh).abs().max() < 1e-4, (i, (got - fresh).abs().max())
        else:                                                     # one flow: forward of the middle centre frame / backward of the last
            sel = 1 if n % 4 == 0 else 5
            got = eng.refine(of, on, oi, select=sel, frame_keys=keys)
            assert (got - fresh[0, sel]).abs().max() < 1e-4, (i, sel, (got - fresh[0, sel]).abs().max())
        rots.append(next(iter(eng._ws.values()))["rot"])
    assert eng.pyramid_slides == 9
    assert rots == ([0, 1, 2, 0, 1, 2, 0, 1, 0, 1, 0, 1] if ring else [0] * 12), rots
