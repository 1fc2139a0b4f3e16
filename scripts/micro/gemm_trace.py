"""VFML_GEMM_TRACE=1 python scripts/micro/gemm_trace.py : per-role wait cycles of the tcgen05 GEMM (built-in debug trace)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
dev = "cuda"
def buf(n, c): return torch.randn((n, (c + 7) // 8 * 8), device=dev).half()[:, :c]
for M, N, K in ((42240, 486, 324), (42240, 324, 486), (42240, 128, 486), (21120, 768, 512), (21120, 512, 768), (21120, 128, 128), (21120, 192, 128)):
    x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
    b = torch.randn(N, device=dev)
    torch.cuda.synchronize()
    print(f"--- {M}x{N}x{K}: warm-up plain / plain / bias+gelu / bias+res+gelu / bias+res+gelu+dw1", file=sys.stderr, flush=True)
    ops.gemm(x, w, o)
    ops.gemm(x, w, o)
    ops.gemm(x, w, o, bias=b, act="gelu")
    ops.gemm(x, w, o, bias=b, res=r, act="gelu")
    ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b))
    torch.cuda.synchronize()
