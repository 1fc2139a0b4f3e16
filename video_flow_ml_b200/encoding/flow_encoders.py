"""Flow encoders / decoders on libvfml kernels, bit-exact with the numpy / torchvision / OpenCV arithmetic of the
reference (encoding/flow_encoders.py: hsv :30-67, gamedev :70-117, motion-vectors-rg8 :120-177, motion-vectors-rgb8
:180-364, torchvision :367-427, factory :430-483, helpers :485-540).

``encode`` accepts a numpy array (uploaded, encoded, downloaded: the reference's signature) or a CUDA tensor (stays on the
device: no host round trip when the flow was just computed there)."""
from abc import ABC, abstractmethod

import numpy as np
import torch

from .. import ops


def _to_device(flow, device="cuda"):
    if isinstance(flow, torch.Tensor):
        return flow.contiguous().float(), True
    return torch.from_numpy(np.ascontiguousarray(flow, dtype=np.float32)).to(device), False


def _u8_to_device(img, device="cuda"):
    if isinstance(img, torch.Tensor):
        return img.contiguous(), True
    return torch.from_numpy(np.ascontiguousarray(img, dtype=np.uint8)).to(device), False


class FlowEncoder(ABC):
    @abstractmethod
    def encode(self, flow, width, height):
        """flow (H, W, 2) -> RGB (H, W, 3) uint8"""


class HSVFlowEncoder(FlowEncoder):
    """Hue = direction, saturation = magnitude / max magnitude, value = 255 (reference :30-67)."""

    def encode(self, flow, width, height):
        t, on_dev = _to_device(flow)
        out = ops.encode_hsv(t)
        return out if on_dev else out.cpu().numpy()


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
        t, on_dev = _u8_to_device(encoded_flow)
        out = ops.decode_mv_rg8(t, self.clamp_range)
        return out if on_dev else out.cpu().numpy()


class MotionVectorsRGB8FlowEncoder(FlowEncoder):
    """Direction in R, G and sqrt(1 - |d|^2) in B: the reference's module-level ``method = 'rgb+'`` (:10, :242-262)."""

    def __init__(self, clamp_range=32.0):
        self.clamp_range = clamp_range

    def encode(self, flow, width, height):
        t, on_dev = _to_device(flow)
        out = ops.encode_mv_rgb8(t, self.clamp_range)
        return out if on_dev else out.cpu().numpy()

    def decode(self, encoded_flow):
        t, on_dev = _u8_to_device(encoded_flow)
        out = ops.decode_mv_rgb8(t, self.clamp_range)
        return out if on_dev else out.cpu().numpy()


class TorchvisionFlowEncoder(FlowEncoder):
    """Colour-wheel format of torchvision.utils.flow_to_image, including the reference's final ``(img * 255).astype(uint8)``
    on the already-uint8 image (:425), which wraps modulo 256.  The fallback encoder argument is kept for signature
    compatibility; the kernel needs no torchvision."""

    def __init__(self, fallback_encoder=None):
        self.fallback_encoder = fallback_encoder or HSVFlowEncoder()

    def encode(self, flow, width, height):
        t, on_dev = _to_device(flow)
        out = ops.encode_torchvision(t, ref_wrap=True)
        return out if on_dev else out.cpu().numpy()


class FlowEncoderFactory:
    _encoders = {
        'hsv': HSVFlowEncoder,
        'gamedev': GamedevFlowEncoder,
        'torchvision': TorchvisionFlowEncoder,
        'motion-vectors-rg8': MotionVectorsRG8FlowEncoder,
        'motion-vectors-rgb8': MotionVectorsRGB8FlowEncoder,
    }

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


def _mv_encoder(clamp_range, format_variant):
    if format_variant.lower() == 'rg8':
        return MotionVectorsRG8FlowEncoder(clamp_range=clamp_range)
    return MotionVectorsRGB8FlowEncoder(clamp_range=clamp_range)       # the reference's default (:502-528)


def encode_motion_vectors(flow, clamp_range=64.0, format_variant='rgb8'):
    h, w = flow.shape[:2]
    return _mv_encoder(clamp_range, format_variant).encode(flow, w, h)


def decode_motion_vectors(encoded_flow, clamp_range=64.0, format_variant='rgb8'):
    return _mv_encoder(clamp_range, format_variant).decode(encoded_flow)
