"""gamedev / motion-vectors-rg8 flow encoders on libvfml kernels, bit-exact with the numpy
reference (encoding/flow_encoders.py:70-177, factory :430-483, helpers :485-540).

``encode`` accepts a numpy array (uploaded, encoded, downloaded: the reference's signature) or a
CUDA tensor (stays on the device: no host round trip when the flow was just computed there)."""
from abc import ABC, abstractmethod

import numpy as np
import torch

from .. import ops


def _to_device(flow, device="cuda"):
    if isinstance(flow, torch.Tensor):
        return flow.contiguous().float(), True
    return torch.from_numpy(np.ascontiguousarray(flow, dtype=np.float32)).to(device), False


class FlowEncoder(ABC):
    @abstractmethod
    def encode(self, flow, width, height):
        """flow (H, W, 2) -> RGB (H, W, 3) uint8"""


class GamedevFlowEncoder(FlowEncoder):
    """R = clip(flow_x / width * scale, +-clamp), G likewise for y, mapped to [0, 255], B = 0."""

    def __init__(self, scale_factor=200.0, clamp_range=20.0):
        self.scale_factor = scale_factor
        self.clamp_range = clamp_range

    def encode(self, flow, width, height):
        t, on_dev = _to_device(flow)
        out = ops.encode_gamedev(t, width, height, self.scale_factor, self.clamp_range)
        return out if on_dev else out.cpu().numpy()


class MotionVectorsRG8FlowEncoder(FlowEncoder):
    def __init__(self, clamp_range=64.0):
        self.clamp_range = clamp_range

    def encode(self, flow, width, height):
        t, on_dev = _to_device(flow)
        out = ops.encode_mv_rg8(t, self.clamp_range)
        return out if on_dev else out.cpu().numpy()

    def decode(self, encoded_flow):
        if isinstance(encoded_flow, torch.Tensor):
            return ops.decode_mv_rg8(encoded_flow.contiguous(), self.clamp_range)
        t = torch.from_numpy(np.ascontiguousarray(encoded_flow, dtype=np.uint8)).cuda()
        return ops.decode_mv_rg8(t, self.clamp_range).cpu().numpy()


class FlowEncoderFactory:
    # hsv / torchvision / motion-vectors-rgb8 of the reference are "next" items
    _encoders = {'gamedev': GamedevFlowEncoder, 'motion-vectors-rg8': MotionVectorsRG8FlowEncoder}

    @classmethod
    def create_encoder(cls, format_name, **kwargs):
        format_name = format_name.lower()
        if format_name not in cls._encoders:
            available_formats = ', '.join(cls._encoders.keys())
            raise ValueError(f"Unsupported format '{format_name}'. Available formats: {available_formats}")
        return cls._encoders[format_name](**kwargs)

    @classmethod
    def get_available_formats(cls):
        return list(cls._encoders.keys())

    @classmethod
    def register_encoder(cls, format_name, encoder_class):
        if not issubclass(encoder_class, FlowEncoder):
            raise ValueError("Encoder class must inherit from FlowEncoder")
        cls._encoders[format_name.lower()] = encoder_class


def encode_flow(flow, width, height, format_name='gamedev'):
    return FlowEncoderFactory.create_encoder(format_name).encode(flow, width, height)


def encode_motion_vectors(flow, clamp_range=64.0, format_variant='rg8'):
    if format_variant.lower() != 'rg8':
        raise NotImplementedError("only the rg8 motion-vector variant is implemented in this build")
    h, w = flow.shape[:2]
    return MotionVectorsRG8FlowEncoder(clamp_range=clamp_range).encode(flow, w, h)


def decode_motion_vectors(encoded_flow, clamp_range=64.0, format_variant='rg8'):
    if format_variant.lower() != 'rg8':
        raise NotImplementedError("only the rg8 motion-vector variant is implemented in this build")
    return MotionVectorsRG8FlowEncoder(clamp_range=clamp_range).decode(encoded_flow)
