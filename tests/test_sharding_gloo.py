"""N > 1 path on CPU: windows sharded over ranks, flows gathered with torch.distributed (gloo, world size 2 and 3)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from video_flow_ml_b200 import sharding


class _FakeProcessor:
    """Stand-in with the VideoFlowProcessor call signature: a deterministic 'flow' of the window's frames."""
    sequence_length = 5

    def compute_optical_flow(self, frames, idx):
        half = self.sequence_length // 2
        ids = [min(max(i, 0), len(frames) - 1) for i in range(idx - half, idx + half + 1)]
        acc = sum(float(k + 1) * frames[i].astype(np.float32).mean(-1) for k, i in enumerate(ids))
        return np.stack([acc, -0.5 * acc + idx], -1)


def _clip(n=7, h=6, w=9):
    r = np.random.default_rng(3)
    return [r.integers(0, 255, (h, w, 3), dtype=np.uint8) for _ in range(n)]


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        clip = sharding.compute_clip_sharded(_FakeProcessor(), _clip(), world, rank)
        assert (clip is None) == (rank != 0)
        if rank == 0:
            np.save(out_path, clip.numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_clip_equals_single_process(tmp_path, world):
    out = str(tmp_path / "clip.npy")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    frames = _clip()
    want = np.stack([_FakeProcessor().compute_optical_flow(frames, i).transpose(2, 0, 1) for i in range(len(frames))])
    assert np.array_equal(np.load(out), want)


def test_shard_ranges_and_tiles():
    for n in (0, 1, 7, 8, 64):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(4, 2, 2)
    tiles = [{'width': 1280, 'height': 1280}, {'width': 1280, 'height': 1280}, {'width': 1280, 'height': 880}]
    items = [sharding.shard_tiles([0, 1], tiles, 4, r) for r in range(4)]
    flat = sorted(sum(items, []))
    assert flat == sorted((f, i) for f in (0, 1) for i in range(3))
    assert all(tiles[it[0][1]]['height'] == 1280 for it in items)      # every rank starts with a large tile
