"""TAAProcessor, TAAComparisonProcessor and apply_taa_effect on the libvfml TAA kernel (reference
effects/taa_processor.py:20-262, :322-425).

History stays on the device as float64 values; the dtype the reference would hold at each point
(float32 on the first frame and in the simple / bilinear modes, float64 after the first
flow+bilateral frame: SURVEY Q10) is tracked so that the arithmetic and the returned numpy dtype
match."""
import numpy as np
import torch

from .. import ops


class TAAProcessor:
    def __init__(self, alpha=0.1, bilateral_sigma_color=25.0, device="cuda"):
        self.alpha = alpha
        self.bilateral_sigma_color = bilateral_sigma_color
        self.device = device
        self.history = {}      # sequence_id -> (device f64 tensor, is_f32)

    def _dev(self, a, dtype):
        if isinstance(a, torch.Tensor):
            return a.to(self.device, dtype).contiguous()
        return torch.from_numpy(np.ascontiguousarray(a)).to(self.device).to(dtype).contiguous()

    def apply_taa(self, current_frame, flow_pixels=None, previous_taa_frame=None, alpha=None, use_flow=True,
                  use_bilateral=True, sequence_id='default', return_device=False):
        alpha = alpha if alpha is not None else self.alpha
        # uint8 frames are uploaded as bytes; anything else goes through float32 like the reference's
        # current_frame.astype(np.float32) (:64), so float frames keep fractions and out-of-range values
        cur = self._dev(current_frame, torch.uint8 if getattr(current_frame, "dtype", None) in (np.uint8, torch.uint8) else torch.float32)
        if previous_taa_frame is not None:
            is32 = getattr(previous_taa_frame, "dtype", None) in (np.float32, torch.float32)
            prev = (self._dev(previous_taa_frame, torch.float64), is32)
        else:
            prev = self.history.get(sequence_id)
        if prev is None:
            state = (ops.taa_init(cur), True)
        else:
            hist, is32 = prev
            if not use_flow or flow_pixels is None:
                state = (ops.taa(cur, None, hist, alpha, self.bilateral_sigma_color, 0, is32), is32)
            elif use_bilateral:
                fl = self._dev(flow_pixels, torch.float32)
                state = (ops.taa(cur, fl, hist, alpha, self.bilateral_sigma_color, 2, is32), False)
            else:
                fl = self._dev(flow_pixels, torch.float32)
                state = (ops.taa(cur, fl, hist, alpha, self.bilateral_sigma_color, 1, True), True)
        self.history[sequence_id] = state
        if return_device:
            return state[0]
        out = state[0].cpu().numpy()
        return out.astype(np.float32) if state[1] else out

    def apply_simple_taa(self, current_frame, previous_taa_frame=None, alpha=None, sequence_id='simple'):
        return self.apply_taa(current_frame, None, previous_taa_frame, alpha, use_flow=False, use_bilateral=False,
                              sequence_id=sequence_id)

    def reset_history(self, sequence_id=None):
        if sequence_id is None:
            self.history.clear()
        else:
            self.history.pop(sequence_id, None)

    def get_history(self, sequence_id='default'):
        st = self.history.get(sequence_id)
        if st is None:
            return None
        out = st[0].cpu().numpy()
        return out.astype(np.float32) if st[1] else out

    def set_alpha(self, alpha):
        if not 0.0 <= alpha <= 1.0:
            raise ValueError("Alpha must be between 0.0 and 1.0")
        self.alpha = alpha


class TAAComparisonProcessor:
    """Flow-based and simple TAA side by side (reference :322-391)."""

    def __init__(self, alpha=0.1):
        self.flow_taa = TAAProcessor(alpha)
        self.simple_taa = TAAProcessor(alpha)

    def apply_comparison(self, current_frame, flow_pixels=None, alpha=None):
        flow_result = self.flow_taa.apply_taa(current_frame=current_frame, flow_pixels=flow_pixels, alpha=alpha,
                                              use_flow=True, use_bilateral=True, sequence_id='flow')
        simple_result = self.simple_taa.apply_simple_taa(current_frame=current_frame, alpha=alpha, sequence_id='simple')
        return flow_result, simple_result

    def reset_history(self):
        self.flow_taa.reset_history()
        self.simple_taa.reset_history()

    def set_alpha(self, alpha):
        self.flow_taa.set_alpha(alpha)
        self.simple_taa.set_alpha(alpha)


def apply_taa_effect(current_frame, flow_pixels=None, previous_taa_frame=None, alpha=0.1, use_flow=True):
    """One-shot convenience call (reference :394-425)."""
    return TAAProcessor(alpha).apply_taa(current_frame=current_frame, flow_pixels=flow_pixels,
                                         previous_taa_frame=previous_taa_frame, alpha=alpha, use_flow=use_flow)
