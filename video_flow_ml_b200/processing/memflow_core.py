"""MemFlowCore on libvfml kernels.  Mirrors processing/memflow_core.py:26-206 of the reference: same constructor,
methods, validation messages and return value ([2, H, W] CPU tensor).

The reference runs every call in a freshly spawned Python process that rebuilds the network, reloads the checkpoint
and creates a new InferenceCore (processing/memflow_inference_isolated.py:22-153); only the arithmetic of that
script is kept: range normalisation (:64-69), InputPadder (:70-71), one ``step`` on the last two frames with
``end=True`` (:73-83), unpad.  The engine is built once in ``load_model``; its memory is cleared per call, which is
what a fresh InferenceCore amounts to.  ``InferenceCore`` below is the stateful mirror for callers that feed a clip
frame by frame (MemFlow/inference/inference_core_skflow.py).
"""
import os

import torch

from ..memflow import MemFlowEngine
from .videoflow_core import _PRECISIONS, pad_amounts


class InferenceCore:
    """inference_core_skflow.py:7-54: ``InferenceCore(network, config)``; ``network`` is a MemFlowEngine."""

    def __init__(self, network, config=None):
        if not isinstance(network, MemFlowEngine):
            raise TypeError("network must be a MemFlowEngine")
        self.model = network
        self.config = config if config is not None else network.cfg
        self.mem_every = network.mem_every
        self.enable_long_term = False
        self.clear_memory()

    def clear_memory(self):
        self.model.clear_memory()

    @property
    def curr_ti(self):
        return self.model.curr_ti

    def step(self, images, end=False, add_pe=False, flow_init=None):
        return self.model.step(images, end=end, add_pe=add_pe, flow_init=flow_init)


class MemFlowCore:
    def __init__(self, device='cuda', model_path='MemFlow_ckpt/MemFlowNet_sintel.pth', stage='sintel', precision=None):
        self.device = self._validate_device(device)
        self.model_path = model_path
        self.stage = stage
        self.model = None
        self.cfg = None
        self.input_padder = None
        self.precision = precision or os.environ.get("VFML_PRECISION", "fp32")
        if self.precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}, got {self.precision}")

    def _validate_device(self, device):
        """memflow_core.py:60-74, minus the CPU fallback: this build has no CPU path."""
        if device == 'auto':
            device = 'cuda'
        if device == 'cuda':
            device = 'cuda:0'
        if not str(device).startswith('cuda'):
            raise RuntimeError("this build runs on CUDA devices only (no CPU fallback)")
        return device

    def load_model(self):
        if self.model is not None:
            print("MemFlow model already loaded")
            return
        if not os.path.exists(self.model_path):
            raise FileNotFoundError(f"MemFlow model not found: {self.model_path}")
        if self.stage not in ('sintel', 'things', 'kitti'):
            raise RuntimeError(f"Failed to load MemFlow model: Unsupported stage: {self.stage}")
        ckpt = torch.load(self.model_path, map_location="cpu")
        self.load_state_dict(ckpt)

    def load_state_dict(self, state_dict, config=None):
        """Build the engine from an in-memory checkpoint (upstream key names, optional 'module.' prefix)."""
        dt = _PRECISIONS[self.precision]
        # encoders stay fp32 in both tiers: the reference's InferenceCore.step runs outside autocast
        # (inference_core_skflow.py:20-54), only its flash read-out is reduced precision
        self.model = MemFlowEngine(state_dict, self.device, dt, config=config)
        self.cfg = self.model.cfg

    def validate_input_tensor(self, tensor):
        """memflow_core.py:124-161"""
        if not isinstance(tensor, torch.Tensor):
            raise TypeError("Input must be a torch.Tensor")
        if len(tensor.shape) != 5:
            raise ValueError(f"Input tensor must be 5D [B, T, C, H, W], got shape: {tensor.shape}")
        B, T, C, H, W = tensor.shape
        if B != 1:
            raise ValueError(f"Batch size must be 1, got: {B}")
        if T < 2:
            raise ValueError(f"Temporal dimension must be at least 2, got: {T}")
        if C != 3:
            raise ValueError(f"Channel dimension must be 3 (RGB), got: {C}")
        if H < 64 or W < 64:
            raise ValueError(f"Spatial dimensions must be at least 64x64, got: {H}x{W}")

    def compute_flow_from_tensor(self, frames_tensor):
        """[1, T, 3, H, W] with values in [0, 255], [0, 1] or [-1, 1] -> [2, H, W] (CPU, like the reference)."""
        if self.model is None:
            raise RuntimeError("Model not loaded. Call load_model() first.")
        self.validate_input_tensor(frames_tensor)
        x = frames_tensor[:, -2:].to(self.device).float()                 # only the last two frames are used (:76)
        max_val = frames_tensor.max().item()                              # over the whole tensor, as :64
        if max_val > 2.0:
            x = 2 * (x / 255.0) - 1.0
        elif max_val > 1.0:
            x = 2 * x - 1.0
        H, W = x.shape[-2:]
        l, r, t, b = pad_amounts(H, W)
        if l or r or t or b:
            x = torch.nn.functional.pad(x, [l, r, t, b, 0, 0], mode='replicate')
        self.model.clear_memory()
        _, flow = self.model.step(x, end=True, crop=(t, l, H, W))
        return flow[0].cpu()

    def get_memory_usage(self):
        allocated = torch.cuda.memory_allocated(self.device) / 1024 ** 3
        reserved = torch.cuda.memory_reserved(self.device) / 1024 ** 3
        return {'allocated_gb': allocated, 'reserved_gb': reserved, 'device': self.device}

    def cleanup(self):
        torch.cuda.empty_cache()
        self.input_padder = None
