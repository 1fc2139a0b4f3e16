"""Host logic of the per-frame feature cache (VideoFlowCore.compute_flow_from_frames) with a stub engine:
which frames are encoded, which are served from the cache, repeated frames of padded windows, eviction.  CPU only."""
import numpy as np
import pytest
import torch

from video_flow_ml_b200.processing.videoflow_core import VideoFlowCore
from video_flow_ml_b200.processing.videoflow_processor import VideoFlowProcessor


class StubEngine:
    """encode_frames = a per-image function; refine returns a digest of its inputs."""

    def __init__(self):
        self.f_calls, self.c_calls = [], []

    def encode_frames(self, x, context=False):
        (self.c_calls if context else self.f_calls).append(x.shape[0])
        m = x.mean(dim=(1, 2, 3)).view(-1, 1, 1, 1)
        if context:
            return m.expand(-1, 128, 2, 3).clone(), (m + 1).expand(-1, 128, 2, 3).clone()
        return (2 * m).expand(-1, 256, 2, 3).clone()

    def encode_new_frames(self, xf, xc):
        return (None if xf is None else self.encode_frames(xf)), (None if xc is None else self.encode_frames(xc, context=True))

    def refine(self, fmaps, net, inp, select=None, crop=None, out=None, frame_keys=None):
        self.last = (fmaps[:, 0, 0, 0].clone(), net[:, 0, 0, 0].clone(), inp[:, 0, 0, 0].clone(), select, crop)
        return torch.zeros(2, crop[2], crop[3])


def test_cache_counts_and_values():
    core = VideoFlowCore("cuda")
    core.model = eng = StubEngine()
    frames = [torch.full((3, 12, 20), float(i)) for i in range(8)]
    uploads = []

    def lazy(i):
        return lambda: (uploads.append(i), frames[i])[1]
    # window centred on 2 (frames 0..4): everything is new
    core.compute_flow_from_frames([lazy(i) for i in range(5)], list(range(5)))
    assert eng.f_calls == [5] and eng.c_calls == [3] and uploads == [0, 1, 2, 3, 4]
    assert eng.last[3] == 3 and eng.last[4] == (2, 2, 12, 20)          # flow index T-2, 'sintel' padding of 12x20 -> 16x24
    assert torch.equal(eng.last[0], 2 * torch.arange(5.0)) and torch.equal(eng.last[1], torch.arange(1.0, 4.0))
    # next window (1..5): one new feature map, one new context (frame 4 was an edge frame before)
    core.compute_flow_from_frames([lazy(i) for i in range(1, 6)], list(range(1, 6)))
    assert eng.f_calls == [5, 1] and eng.c_calls == [3, 1] and uploads[5:] == [5, 4]
    assert torch.equal(eng.last[0], 2 * torch.arange(1.0, 6.0)) and torch.equal(eng.last[2], torch.arange(3.0, 6.0))
    # clip start: window of frame 0 = [0, 0, 0, 1, 2] -> frame 0 needs a context now, encoded ONCE for its three copies
    core.compute_flow_from_frames([lazy(i) for i in (0, 0, 0, 1, 2)], [0, 0, 0, 1, 2])
    assert eng.f_calls == [5, 1] and eng.c_calls == [3, 1, 1]
    assert core.cache_stats == {"fnet_runs": 6, "cnet_runs": 5, "fnet_hits": 9, "cnet_hits": 3}
    # uncached keys (None) are always recomputed, and repeated None frames are not merged
    core.compute_flow_from_frames([frames[7]] * 3, [None, None, None])
    assert eng.f_calls[-1] == 3 and eng.c_calls[-1] == 1
    # eviction
    core.feature_cache_frames = 2
    core.compute_flow_from_frames([lazy(i) for i in range(3)], [10, 11, 12])
    assert list(core._feat) == [11, 12]
    with pytest.raises(ValueError):
        core.compute_flow_from_frames([frames[0]] * 2, [0, 1])
    core.model = None
    with pytest.raises(RuntimeError):
        core.compute_flow_from_frames([frames[0]] * 3, [0, 1, 2])


def test_frame_keys_notice_identity_and_content():
    a = np.zeros((16, 24, 3), np.uint8)
    b = a.copy()
    k = VideoFlowProcessor._frame_key
    assert k(a) == k(a) and k(a) != k(b)              # different arrays
    ka = k(a)
    a[:] = 7
    assert k(a) != ka                                 # same array, rewritten in place
    assert k(a[:8]) != k(a)


def test_feature_arena_slots_capacity_and_eviction():
    """Arena bookkeeping of the feature cache (CUDA path of _ensure_features; exercised here with CPU tensors): capacity from the
    frame and byte limits, never below two windows; a full arena hands out the slot of its least recently used frame that is not
    part of the current window; caches smaller than two windows are not pooled."""
    core = VideoFlowCore("cuda")
    fm = torch.zeros(256, 4, 6)
    core.feature_cache_frames = 12
    core.feature_cache_bytes = 7 * 2 * fm.numel() * 4            # room for 7 entries: below two windows of T = 5 -> 10 slots
    keep = []
    for k in range(10):
        arena, slot = core._arena_slot(fm, 5, keep)
        assert slot == k and arena.cap == 10
        core._feat[k] = {"fmap": arena.put_fmap(slot, fm + k), "ctx": None, "arena": arena, "slot": slot}
    assert not arena.free and arena.fmap.shape == (10, 256, 4, 6) and arena.net is None
    cur = [core._feat[0], core._feat[1]]                          # the current window holds the two oldest frames
    a2, slot = core._arena_slot(fm, 5, cur)
    assert a2 is arena and slot == 2 and list(core._feat) == [0, 1, 3, 4, 5, 6, 7, 8, 9]
    net, inp = arena.put_ctx(slot, torch.ones(128, 4, 6), 2 * torch.ones(128, 4, 6))
    assert arena.net.shape == (10, 128, 4, 6) and float(net.sum()) == 128 * 24 and float(inp[0, 0, 0]) == 2.0
    assert float(core._feat[3]["fmap"][0, 0, 0]) == 3.0          # other slots untouched
    assert core._arena_slot(torch.zeros(256, 8, 8), 5, [])[0] is not arena        # another shape, another arena
    core.feature_cache_frames = 9                                 # < 2 windows: not pooled
    assert core._arena_slot(fm, 5, []) is None
    core.clear_feature_cache()
    assert not core._arenas and not core._feat
