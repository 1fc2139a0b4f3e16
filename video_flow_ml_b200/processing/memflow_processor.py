"""MemFlowProcessor / MemFlowInference mirrors (processing/memflow_processor.py:19-256, memflow_inference.py:20-109):
frame validation, window preparation and the (no-op) tile interface, delegating to MemFlowCore."""
import numpy as np
import torch

from .memflow_core import MemFlowCore


class MemFlowProcessor:
    def __init__(self, device='cuda', model_path='MemFlow_ckpt/MemFlowNet_sintel.pth', stage='sintel', sequence_length=3,
                 precision=None):
        self.device = device
        self.model_path = model_path
        self.stage = stage
        self.sequence_length = max(2, sequence_length)
        self.core_engine = MemFlowCore(device=device, model_path=model_path, stage=stage, precision=precision)

    def load_model(self):
        self.core_engine.load_model()

    def validate_frame_sequence(self, frames):
        """memflow_processor.py:57-91"""
        if not frames:
            raise ValueError("Frame sequence is empty")
        if len(frames) < 2:
            raise ValueError(f"Need at least 2 frames for optical flow, got {len(frames)}")
        first = frames[0]
        if len(first.shape) != 3:
            raise ValueError(f"Frames must be 3D (H, W, C), got shape: {first.shape}")
        height, width, channels = first.shape
        if channels != 3:
            raise ValueError(f"Frames must have 3 channels (RGB), got: {channels}")
        if height < 64 or width < 64:
            raise ValueError(f"Frame dimensions must be at least 64x64, got: {height}x{width}")
        for i, frame in enumerate(frames[1:], 1):
            if frame.shape != first.shape:
                raise ValueError(f"Frame {i} shape {frame.shape} differs from first frame {first.shape}")
        return True

    def prepare_frame_sequence(self, frames, frame_idx):
        """The `sequence_length` frames ending at frame_idx (front-padded by repetition), uint8-clipped, as float
        [1, T, 3, H, W] holding 0..255.  memflow_processor.py:93-135"""
        self.validate_frame_sequence(frames)
        end_idx = frame_idx + 1
        start_idx = max(0, end_idx - self.sequence_length)
        seq = list(frames[start_idx:end_idx])
        while len(seq) < self.sequence_length:
            seq.insert(0, seq[0])
        ts = []
        for frame in seq:
            if frame.dtype != np.uint8:
                frame = np.clip(frame, 0, 255).astype(np.uint8)
            ts.append(torch.from_numpy(frame).permute(2, 0, 1).float())
        return torch.stack(ts, dim=0).unsqueeze(0)

    def compute_optical_flow(self, frames, frame_idx):
        """-> np.float32 [H, W, 2]: flow from frame_idx - 1 to frame_idx.  memflow_processor.py:137-157"""
        x = self.prepare_frame_sequence(frames, frame_idx)
        flow = self.core_engine.compute_flow_from_tensor(x)
        return flow.permute(1, 2, 0).numpy()

    def compute_optical_flow_with_progress(self, frames, frame_idx, tile_pbar=None):
        if tile_pbar is not None:
            tile_pbar.set_description("MemFlow processing")
            tile_pbar.update(1)
        return self.compute_optical_flow(frames, frame_idx)

    def calculate_tile_grid(self, width, height, tile_size=1280):
        """MemFlow does not tile: one tile covering the frame.  memflow_processor.py:180-200"""
        return width, height, 1, 1, [{'x': 0, 'y': 0, 'width': width, 'height': height, 'tile_idx': 0}]

    def extract_tile(self, frame, tile_info):
        return frame

    def compute_optical_flow_tiled(self, frames, frame_idx, tile_pbar=None, overall_pbar=None):
        for bar, text in ((tile_pbar, "MemFlow full-frame"), (overall_pbar, "MemFlow processing")):
            if bar is not None:
                bar.set_description(text)
                bar.reset(total=1)
                bar.update(1)
        return self.compute_optical_flow(frames, frame_idx)

    def get_core_engine(self):
        return self.core_engine

    def get_memory_usage(self):
        return self.core_engine.get_memory_usage()

    def cleanup(self):
        self.core_engine.cleanup()


class MemFlowInference:
    """Compatibility layer, memflow_inference.py:20-109: delegates everything to MemFlowProcessor."""

    def __init__(self, device='cuda', model_path='MemFlow_ckpt/MemFlowNet_sintel.pth', stage='sintel', sequence_length=3,
                 precision=None):
        self.device, self.model_path, self.stage, self.sequence_length = device, model_path, stage, sequence_length
        self.processor = MemFlowProcessor(device=device, model_path=model_path, stage=stage, sequence_length=sequence_length,
                                          precision=precision)
        self.model = None
        self.cfg = None

    def load_model(self):
        self.processor.load_model()
        core = self.processor.get_core_engine()
        self.model, self.cfg = core.model, core.cfg

    def __getattr__(self, name):
        # prepare_frame_sequence, compute_optical_flow[_with_progress|_tiled], calculate_tile_grid, extract_tile,
        # get_core_engine, get_memory_usage, cleanup: plain delegation, as in the reference
        if name in ("prepare_frame_sequence", "compute_optical_flow", "compute_optical_flow_with_progress",
                    "compute_optical_flow_tiled", "calculate_tile_grid", "extract_tile", "get_core_engine",
                    "get_memory_usage", "cleanup"):
            return getattr(self.processor, name)
        raise AttributeError(name)

    def get_processor(self):
        return self.processor
