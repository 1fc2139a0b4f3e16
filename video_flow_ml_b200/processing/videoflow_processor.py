"""VideoFlowProcessor on libvfml kernels.  Mirrors processing/videoflow_processor.py:20-361 of the
reference (window choice, tile grid, progress-bar calls, validation, return types)."""
import collections

import numpy as np
import torch

from .videoflow_core import VideoFlowCore


class _NoEvent:
    def record(self):
        pass

    def synchronize(self):
        pass


class VideoFlowProcessor:
    def __init__(self, device, fast_mode=False, tile_mode=False, sequence_length=5,
                 dataset='sintel', architecture='mof', variant='standard', precision=None):
        self.device = device
        self.fast_mode = fast_mode
        self.tile_mode = tile_mode
        self.sequence_length = sequence_length
        self.dataset = dataset
        self.architecture = architecture
        self.variant = variant
        self.core = VideoFlowCore(device, fast_mode, dataset, architecture, variant, precision=precision)
        # sliding windows over one clip share T-1 frames: their uploads and encoder outputs are cached by frame identity
        # (array address, layout and a strided content sample).  set_feature_cache(False) restores one self-contained
        # computation per call, for callers that rewrite frame arrays in place between calls.
        self.feature_cache = True
        self._dev_frames = collections.OrderedDict()      # frame key -> device tensor [H, W, 3]
        self.device_frame_capacity = None                 # frames kept on the device (default: two windows' worth)

    def load_model(self):
        model_path = self.core.load_model()
        print(f"VideoFlow model loaded successfully from: {model_path}")

    # ------------------------------------------------------------------ host logic
    def calculate_tile_grid(self, width, height, tile_size=1280):
        """Disjoint fixed-size tiles, ragged at the right / bottom edges (reference :73-110)."""
        cols = -(-width // tile_size)
        rows = -(-height // tile_size)
        tiles_info = [{'x': c * tile_size, 'y': r * tile_size,
                       'width': min(tile_size, width - c * tile_size), 'height': min(tile_size, height - r * tile_size),
                       'col': c, 'row': r} for r in range(rows) for c in range(cols)]
        return tile_size, tile_size, cols, rows, tiles_info

    def extract_tile(self, frame, tile_info):
        x, y, w, h = tile_info['x'], tile_info['y'], tile_info['width'], tile_info['height']
        return frame[y:y + h, x:x + w]

    def window_indices(self, n_frames, frame_idx):
        """Centred window clipped to the clip; short windows repeat the first frame when clipped at the
        start, otherwise the last (reference :134-150)."""
        half = self.sequence_length // 2
        start = max(0, frame_idx - half)
        ids = list(range(start, min(n_frames, frame_idx + half + 1)))
        while len(ids) < self.sequence_length:
            if start == 0:
                ids.insert(0, ids[0])
            else:
                ids.append(ids[-1])
        return ids[:self.sequence_length]

    def prepare_frame_sequence(self, frames, frame_idx):
        """-> [1, T, 3, H, W] fp32 on the device; uint8 frames are scaled to [0, 1], float frames pass
        through (reference :152-162)."""
        ids = self.window_indices(len(frames), frame_idx)
        seq = [frames[i] for i in ids]
        if all(f.dtype == np.uint8 for f in seq) and len({f.shape for f in seq}) == 1:
            # one u8 upload from a pinned staging buffer (two buffers alternate; a buffer is rewritten only after the
            # copy that last read it has finished), conversion on the device (same fp32 division as the reference's numpy one)
            u8 = self._stage_u8(seq).to(self.device, non_blocking=True)
            self._stage_events[self._stage_i].record()
            return (u8.permute(0, 3, 1, 2).float() / 255.0).unsqueeze(0)
        tensors = [torch.from_numpy((f.astype(np.float32) / 255.0) if f.dtype == np.uint8 else f.astype(np.float32)).permute(2, 0, 1)
                   for f in seq]
        return torch.stack(tensors).unsqueeze(0).to(self.device)

    def _stage_u8(self, seq):
        """Copies the window's frames into one of two pinned [T, H, W, 3] uint8 staging tensors."""
        shape = (len(seq),) + tuple(seq[0].shape)
        st = getattr(self, "_stage", None)
        if st is None or tuple(st[0].shape) != shape:
            pin = str(self.device).startswith("cuda") and torch.cuda.is_available()
            self._stage = st = [torch.empty(shape, dtype=torch.uint8, pin_memory=pin) for _ in range(2)]
            self._stage_np = [t.numpy() for t in st]
            self._stage_events = [torch.cuda.Event() for _ in range(2)] if pin else [_NoEvent(), _NoEvent()]
            self._stage_i = 0
        self._stage_i ^= 1
        self._stage_events[self._stage_i].synchronize()
        dst = self._stage_np[self._stage_i]
        for i, f in enumerate(seq):
            np.copyto(dst[i], f)
        return st[self._stage_i]

    # ------------------------------------------------------------------ per-frame cache
    def set_feature_cache(self, enabled):
        self.feature_cache = bool(enabled)
        if not enabled:
            self._dev_frames.clear()
            self.core.clear_feature_cache()

    @staticmethod
    def _frame_key(f):
        """Identity of a frame array: address, layout, and ~4096 strided ELEMENTS of content (an in-place edit of the array
        is noticed unless it misses every sampled element; callers that do that should disable the cache).  Whole elements at
        an odd stride: a byte stride that is a multiple of the item size would only ever see, e.g., the always-zero low
        mantissa byte of integer-valued float32 frames, and an even one only one colour channel."""
        flat = f.reshape(-1) if f.flags.c_contiguous else np.ascontiguousarray(f).reshape(-1)
        step = max(1, flat.size // 4096) | 1
        return (f.__array_interface__['data'][0], f.shape, f.dtype.str, f.strides, hash(flat[::step].tobytes()))

    def _device_frame(self, f, key):
        """The frame on the device ([H, W, 3], uint8 or float32), uploaded once per key through pinned staging."""
        t = self._dev_frames.get(key)
        if t is not None:
            self._dev_frames.move_to_end(key)
            if t.is_cuda:
                t.record_stream(torch.cuda.current_stream(t.device))      # uploaded on the prefetch stream, read on this one (or vice versa)
            return t
        if f.dtype == np.uint8:
            ring = getattr(self, "_pin_ring", None)
            if ring is None or tuple(ring[0][0].shape) != f.shape:
                pin = str(self.device).startswith("cuda") and torch.cuda.is_available()
                ring = self._pin_ring = [(torch.empty(f.shape, dtype=torch.uint8, pin_memory=pin),
                                          torch.cuda.Event() if pin else _NoEvent()) for _ in range(8)]
                self._pin_i = 0
            self._pin_i = (self._pin_i + 1) % len(ring)
            buf, ev = ring[self._pin_i]
            ev.synchronize()                   # the copy that last read this staging buffer has finished
            np.copyto(buf.numpy(), f)
            t = buf.to(self.device, non_blocking=True)
            ev.record()
        else:
            t = torch.from_numpy(np.ascontiguousarray(f, dtype=np.float32)).to(self.device)
        self._dev_frames[key] = t
        while len(self._dev_frames) > (self.device_frame_capacity or 2 * max(self.sequence_length, 5)):
            self._dev_frames.popitem(last=False)
        return t

    @staticmethod
    def _chw(t, rect=None):
        """Device frame [H, W, 3] (optionally a tile of it) -> [3, h, w] fp32 as prepare_frame_sequence would make it."""
        if rect is not None:
            x, y, w, h = rect
            t = t[y:y + h, x:x + w]
        t = t.permute(2, 0, 1)
        return t.float() / 255.0 if t.dtype == torch.uint8 else t.float()

    def _cached_window(self, frames, frame_idx):
        """(frames on demand, keys) of the window for VideoFlowCore.compute_flow_from_frames, or None when the frames
        are not plain same-shape arrays (then the uncached tensor path is used)."""
        ids = self.window_indices(len(frames), frame_idx)
        seq = [frames[i] for i in ids]
        if not (self.feature_cache and self.sequence_length >= 3 and all(isinstance(f, np.ndarray) and f.ndim == 3 for f in seq)
                and len({(f.shape, f.dtype) for f in seq}) == 1):
            return None
        keys = [self._frame_key(f) for f in seq]
        return seq, keys

    def _lazy_window(self, cw, rect=None):
        """(frames on demand, keys) for VideoFlowCore.compute_flow_from_frames; rect = (x, y, w, h) selects a tile."""
        seq, keys = cw
        lazy = [(lambda f=f, k=k: self._chw(self._device_frame(f, k), rect)) for f, k in zip(seq, keys)]
        return lazy, (keys if rect is None else [k + rect for k in keys])

    def _next_window(self, frames, frame_idx):
        """The window a sliding pass asks for next (frame_idx + 1), for the core's encoder prefetch."""
        if frame_idx + 1 >= len(frames):
            return None
        cw = self._cached_window(frames, frame_idx + 1)
        return None if cw is None else self._lazy_window(cw)

    # ------------------------------------------------------------------ inference
    def compute_optical_flow(self, frames, frame_idx, return_device=False):
        """-> np.float32 [H, W, 2] like the reference (:164-187).  return_device=True keeps the [H, W, 2] flow on the GPU
        (a view of the kernel's [2, H, W] output) for the device-resident post-pass: encoding.encode_flow and
        effects.TAAProcessor accept CUDA tensors, so flow -> encode -> TAA needs no host round trip."""
        if not self.core.is_model_loaded():
            raise RuntimeError("Model not loaded. Call load_model() first.")
        cw = self._cached_window(frames, frame_idx)
        if cw is not None:
            flow = self.core.compute_flow_from_frames(*self._lazy_window(cw), next_window=self._next_window(frames, frame_idx))
        else:
            flow = self.core.compute_flow_from_tensor(self.prepare_frame_sequence(frames, frame_idx))
        if return_device:
            return flow.permute(1, 2, 0)
        return flow.permute(1, 2, 0).cpu().numpy()

    def compute_optical_flow_with_progress(self, frames, frame_idx, tile_pbar=None):
        if not self.core.is_model_loaded():
            raise RuntimeError("Model not loaded. Call load_model() first.")
        if tile_pbar is not None:
            tile_pbar.set_description("Preparing frames")
            tile_pbar.reset()
        batch = self.prepare_frame_sequence(frames, frame_idx)
        if tile_pbar is not None:
            tile_pbar.set_description("Running VideoFlow")
            tile_pbar.update(2)
        flow = self.core.compute_flow_from_tensor(batch)
        if tile_pbar is not None:
            tile_pbar.set_description("Processing output")
            tile_pbar.update(1)
        flow_np = flow.permute(1, 2, 0).cpu().numpy()
        if tile_pbar is not None:
            tile_pbar.set_description("Completed")
            tile_pbar.update(1)
        return flow_np

    def compute_optical_flow_tiled(self, frames, frame_idx, tile_pbar=None, overall_pbar=None, return_device=False):
        """Disjoint 1280x1280 tiles, each an independent clip, pasted without blending (reference :231-283).
        Tiles are cropped from the window's frames only and each tile's flow is written by the upsampling
        kernel straight into the full-frame device buffer (no per-tile D2H, no host paste)."""
        if not self.tile_mode:
            return self.compute_optical_flow(frames, frame_idx, return_device=return_device)
        if not self.core.is_model_loaded():
            raise RuntimeError("Model not loaded. Call load_model() first.")
        height, width = frames[frame_idx].shape[:2]
        _, _, _, _, tiles_info = self.calculate_tile_grid(width, height)
        full = torch.zeros((2, height, width), dtype=torch.float32, device=self.device)
        cw = self._cached_window(frames, frame_idx)
        window = self.prepare_frame_sequence(frames, frame_idx) if cw is None else None
        nxt = self._cached_window(frames, frame_idx + 1) if (cw is not None and frame_idx + 1 < len(frames)) else None
        for i, ti in enumerate(tiles_info):
            if overall_pbar is not None:
                overall_pbar.set_description(f"Tile {i+1}/{len(tiles_info)} ({ti['width']}x{ti['height']})")
            if tile_pbar is not None:
                tile_pbar.set_description("Running VideoFlow")
                tile_pbar.reset()
            self._compute_tile(cw, window, ti, full, nxt)
            if tile_pbar is not None:
                tile_pbar.set_description("Completed")
                tile_pbar.update(4)
            if overall_pbar is not None:
                overall_pbar.update(1)
        if return_device:
            return full.permute(1, 2, 0)
        return full.permute(1, 2, 0).cpu().numpy()

    def _compute_tile(self, cw, window, ti, full, nxt=None):
        """One tile of one output frame, written by the upsampling kernel into its place in `full` [2, H, W].
        nxt: the cached window of the next output frame (its new frame's tile is encoded beside this tile's iterations)."""
        x, y, w, h = ti['x'], ti['y'], ti['width'], ti['height']
        if cw is not None:          # per-(frame, tile) encoder cache; the frame itself is uploaded once for all its tiles
            seq, keys = cw
            rect = (x, y, w, h)
            self.core.compute_flow_from_frames(*self._lazy_window(cw, rect), out=full[:, y:y + h, x:x + w],
                                               next_window=None if nxt is None else self._lazy_window(nxt, rect))
        else:
            tile = window[..., y:y + h, x:x + w].contiguous()
            self.core.compute_flow_from_tensor(tile, out=full[:, y:y + h, x:x + w])

    def compute_tile_into(self, frames, frame_idx, tile_info, full):
        """Sharded tile mode (sharding.shard_tile_chunks): compute ONE tile of frame `frame_idx` into `full` [2, H, W]."""
        cw = self._cached_window(frames, frame_idx)
        window = self.prepare_frame_sequence(frames, frame_idx) if cw is None else None
        nxt = self._cached_window(frames, frame_idx + 1) if (cw is not None and frame_idx + 1 < len(frames)) else None
        self._compute_tile(cw, window, tile_info, full, nxt)

    # ------------------------------------------------------------------ misc API
    def is_model_loaded(self):
        return self.core.is_model_loaded()

    def get_model_info(self):
        info = self.core.get_model_info()
        if info["status"] == "loaded":
            info.update({"tile_mode": self.tile_mode, "sequence_length": self.sequence_length,
                         "processor_type": "VideoFlowProcessor"})
        return info

    def get_memory_usage(self):
        return self.core.get_memory_usage()

    def validate_frames(self, frames, frame_idx):
        if not isinstance(frames, list):
            raise ValueError("Frames must be a list of numpy arrays")
        if len(frames) == 0:
            raise ValueError("Frames list cannot be empty")
        if frame_idx < 0 or frame_idx >= len(frames):
            raise ValueError(f"Frame index {frame_idx} out of range [0, {len(frames)-1}]")
        f = frames[0]
        if not isinstance(f, np.ndarray):
            raise ValueError("Frames must be numpy arrays")
        if len(f.shape) != 3:
            raise ValueError(f"Frames must be 3D arrays [H,W,C], got shape {f.shape}")
        if f.shape[2] != 3:
            raise ValueError(f"Frames must have 3 color channels, got {f.shape[2]}")
        if f.dtype not in [np.uint8, np.float32, np.float64]:
            raise ValueError(f"Unsupported frame dtype: {f.dtype}")
        if f.dtype != np.uint8 and (f.max() > 1.0 or f.min() < 0.0):
            if not (f.max() <= 255.0 and f.min() >= 0.0):
                raise ValueError("Float frames must be in range [0.0, 1.0] or [0.0, 255.0]")

    def set_tile_mode(self, enabled):
        self.tile_mode = enabled

    def set_sequence_length(self, length):
        if length < 1 or length > 10:
            raise ValueError("Sequence length must be between 1 and 10")
        self.sequence_length = length
