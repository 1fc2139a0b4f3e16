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
