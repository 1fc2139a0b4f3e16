"""Only the tensor-core depthwise kernels at config C1 shapes (for ncu captures and A/B timing).
python scripts/micro/dw_only.py [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops

h, w = 55, 128
P = h * w
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
for name, n_img, C, k in [("gru.dw7", 3, 512, 7), ("convc1.dw15", 6, 324, 15), ("conv.dw15", 3, 400, 15), ("fhead.dw15", 3, 128, 15)]:
    M = n_img * P
    cp = (C + 7) // 8 * 8
    x = torch.randn((M, cp), device="cuda").half()[:, :C]
    y = torch.empty((M, cp), device="cuda", dtype=torch.float16)[:, :C]
    b = torch.randn(C, device="cuda")
    wp = ops.dwconv_pack_weights(torch.randn((k * k, C), device="cuda") / k, k)
    fn = lambda: ops.dwconv_tc_res_gelu(x, y, n_img, h, w, wp, b, k)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record()
    torch.cuda.synchronize()
    print(f"{name:14s} C={C:4d} n_img={n_img} {e0.elapsed_time(e1) / reps * 1000:8.1f} us   checksum {y.float().abs().mean().item():.6f}")
