"""Sharding of a clip's sliding windows (and 4K tiles) over the GPUs of one node.

Each output frame's window is a pure function of its T input frames (reference
processing/videoflow_processor.py:164-187: no state crosses windows), so the frame indices are split
into contiguous ranges, one per rank (one process per GPU), with NO data-path collective: every rank
reads its range plus a floor(T/2)-frame halo from the host-resident clip.  The only communication is
the gather of the finished flow fields at the end (torch.distributed: NCCL on GPUs, gloo in the CPU
tests).  Tiles of one frame are independent in the same way (reference :231-283, no blending) and are
dealt out as (frame, tile) work items sorted by size.
"""
import torch
import torch.distributed as dist


def shard_range(n_items, world_size, rank):
    """Contiguous [start, stop) of rank `rank`; sizes differ by at most one, earlier ranks get the larger shards."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank {rank} / world size {world_size}")
    base, extra = divmod(n_items, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_tiles(frame_ids, tiles_info, world_size, rank):
    """(frame, tile) work items for `rank`: largest tiles first, dealt round-robin so ragged edge tiles balance."""
    items = [(f, i) for f in frame_ids for i in range(len(tiles_info))]
    items.sort(key=lambda it: (-tiles_info[it[1]]['width'] * tiles_info[it[1]]['height'], it[0], it[1]))
    return items[rank::world_size]


def shard_tile_chunks(n_frames, tiles_info, world_size, rank, chunk=4):
    """Tile mode over one clip: work units are (tile, contiguous frame range) chunks, so that consecutive frames of the
    SAME tile stay on one rank and its per-(frame, tile) encoder cache keeps working (a (frame, tile) item costs T + (T-2)
    encoder passes cold and 1 + 1 behind its predecessor).  Chunks are dealt largest-first to the least-loaded rank
    (ragged edge tiles weigh less).  -> this rank's list of (tile_index, frame_start, frame_stop), and every rank's load."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank {rank} / world size {world_size}")
    units = []
    for ti, t in enumerate(tiles_info):
        for a in range(0, n_frames, chunk):
            b = min(n_frames, a + chunk)
            units.append((t['width'] * t['height'] * (b - a), ti, a, b))
    units.sort(key=lambda u: (-u[0], u[1], u[2]))
    load = [0] * world_size
    mine = []
    for cost, ti, a, b in units:
        r = min(range(world_size), key=lambda i: (load[i], i))
        load[r] += cost
        if r == rank:
            mine.append((ti, a, b))
    mine.sort()
    return mine, load


def gather_flows(local, n_total, world_size, rank, dst=0, group=None):
    """local: [n_local, 2, H, W] flows of this rank's contiguous shard (shard_range order).
    Returns the [n_total, 2, H, W] clip on `dst`, None elsewhere.  Shards may differ in length by one, so they
    are padded to the longest for the fixed-size gather and trimmed on arrival."""
    if world_size == 1:
        return local
    longest = shard_range(n_total, world_size, 0)[1]
    buf = local
    if local.shape[0] < longest:
        buf = torch.cat([local, local.new_zeros((longest - local.shape[0],) + tuple(local.shape[1:]))])
    parts = [torch.empty_like(buf) for _ in range(world_size)] if rank == dst else None
    dist.gather(buf.contiguous(), parts, dst=dst, group=group)
    if rank != dst:
        return None
    out = []
    for r, part in enumerate(parts):
        a, b = shard_range(n_total, world_size, r)
        out.append(part[: b - a])
    return torch.cat(out)


def compute_clip_sharded(processor, frames, world_size, rank, device=None, dst=0):
    """Every rank computes the flow of its contiguous range of frames with `processor`
    (VideoFlowProcessor API: compute_optical_flow(frames, idx) -> np [H, W, 2]); rank `dst` returns the
    whole clip [n, 2, H, W], the others None."""
    a, b = shard_range(len(frames), world_size, rank)
    H, W = frames[0].shape[:2]
    local = torch.empty((b - a, 2, H, W), dtype=torch.float32, device=device)
    for i in range(a, b):
        local[i - a] = torch.as_tensor(processor.compute_optical_flow(frames, i)).permute(2, 0, 1)
    return gather_flows(local, len(frames), world_size, rank, dst)
