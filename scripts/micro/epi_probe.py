"""Decode the epilogue's row / column mapping: D[m, n] = m and D[m, n] = n through the tensor-core GEMM."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
for (M, N, K) in ((128, 128, 64), (128, 64, 64), (300, 176, 128)):
    a = torch.zeros((M, K), device="cuda", dtype=torch.float16); b = torch.zeros((N, K), device="cuda", dtype=torch.float16)
    for name in ("row", "col"):
        a.zero_(); b.zero_()
        if name == "row":
            a[:, 0] = torch.arange(M, device="cuda").half(); b[:, 0] = 1
        else:
            a[:, 0] = 1; b[:, 0] = torch.arange(N, device="cuda").half()
        for odt in (torch.float16, torch.float32):
            out = torch.full((M, N), -1.0, device="cuda", dtype=odt)
            ops.gemm(a, b, out)
            torch.cuda.synchronize()
            want = (torch.arange(M, device="cuda").view(-1, 1).expand(M, N) if name == "row" else torch.arange(N, device="cuda").view(1, -1).expand(M, N)).float()
            bad = (out.float() != want)
            print(f"{M}x{N}x{K} D={name} out={odt}: {int(bad.sum())} wrong of {M * N}")
            if bad.any():
                print(" got row 0, cols 0..40:", [int(v) for v in out[0, 0:40].float().tolist()])
                print(" got row 1, cols 0..40:", [int(v) for v in out[1, 0:40].float().tolist()])
                print(" got col 0, rows 0..40:", [int(v) for v in out[0:40, 0].float().tolist()])
