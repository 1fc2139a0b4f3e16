"""VFML_GEMM_TRACE=1 python scripts/micro/gemm_trace.py : per-role wait cycles of the tcgen05 GEMM (debug build-in trace)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
dev = "cuda"
def buf(n, c): return torch.randn((n, (c + 7) // 8 * 8), device=dev).half()[:, :c]
for M, N, K in ((42240, 486, 324), (42240, 486, 64), (42240, 256, 324), (21120, 768, 512), (21120, 128, 128)):
    x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
    b = torch.randn(N, device=dev)
    for _ in range(3): ops.gemm(x, w, o)
    ops.gemm(x, w, o, bias=b, act="gelu")
    ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b))
    torch.cuda.synchronize()
