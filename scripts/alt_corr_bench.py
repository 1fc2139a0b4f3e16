"""Times the volume-free lookup (vfml_alt_corr_forward, 4 levels) at config C1 (3 frames, 55 x 128, C = 256), graph-replayed."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from video_flow_ml_b200 import ops

torch.manual_seed(0)
dev = "cuda"
B, h, w, C = 3, 55, 128, 256
for dt in (torch.float32, torch.float16):
    f1 = torch.randn(B, h, w, C, device=dev).to(dt)
    f2s, f = [], torch.randn(B, C, h, w, device=dev)
    for l in range(4):
        f2s.append(f.permute(0, 2, 3, 1).contiguous().to(dt))
        f = torch.nn.functional.avg_pool2d(f, 2, stride=2)
    ys, xs = torch.meshgrid(torch.arange(h, device=dev), torch.arange(w, device=dev), indexing="ij")
    base = torch.stack([xs, ys], -1).float().view(1, 1, h, w, 2).expand(B, 1, h, w, 2)
    sets = [(base + 4 * torch.randn(B, 1, h, w, 2, device=dev)).contiguous() for _ in range(4)]
    sets = [[(c / 2 ** l).contiguous() for l in range(4)] for c in sets]
    for c in sets:
        for l in range(4):
            ops.alt_corr_forward(f1, f2s[l], c[l], 4)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    reps = 8
    with torch.cuda.graph(g):
        for i in range(reps):
            for l in range(4):
                ops.alt_corr_forward(f1, f2s[l], sets[i % 4][l], 4)
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    flop = 2.0 * B * h * w * 4 * 100 * C
    print(f"alt_corr 4 levels {dt}: {us:8.1f} us per direction ({flop / us / 1e6:.2f} TFLOP/s useful)")
