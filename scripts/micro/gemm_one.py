"""One large GEMM shape, a few launches (for ncu source-level sampling): python scripts/micro/gemm_one.py [plain|gelu|res]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from video_flow_ml_b200 import ops
mode = sys.argv[1] if len(sys.argv) > 1 else "plain"
dev = "cuda"
def buf(n, c): return torch.randn((n, (c + 15) // 16 * 16), device=dev).half()[:, :c]
M, N, K = 42240, 486, 324
x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
b = torch.randn(N, device=dev)
for _ in range(6):
    if mode == "plain": ops.gemm(x, w, o, weights=True)
    elif mode == "gelu": ops.gemm(x, w, o, bias=b, act="gelu", weights=True)
    else: ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b), weights=True)
torch.cuda.synchronize()
