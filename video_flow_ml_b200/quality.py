"""Flow quality map on a libvfml kernel: mirrors ``generate_quality_frame_gpu(frame1, frame2, flow, device,
good_quality_threshold)`` of the reference (correction_worker.py:175-208) -- numpy u8 frames and an fp32 flow
in, numpy u8 [H, W, 3] out; CUDA tensors in -> CUDA tensor out."""
import numpy as np
import torch

from . import ops


def generate_quality_frame_gpu(frame1, frame2, flow, device="cuda", good_quality_threshold=0.7):
    on_dev = all(torch.is_tensor(t) and t.is_cuda for t in (frame1, frame2, flow))
    up8 = lambda f: f.contiguous() if torch.is_tensor(f) else torch.from_numpy(np.ascontiguousarray(f, dtype=np.uint8)).to(device)
    fl = flow.contiguous().float() if torch.is_tensor(flow) else torch.from_numpy(np.ascontiguousarray(flow, dtype=np.float32)).to(device)
    out = ops.quality_frame(up8(frame1), up8(frame2), fl.to(device), good_quality_threshold)
    return out if on_dev else out.cpu().numpy()
