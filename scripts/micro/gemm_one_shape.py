"""One update-block layer (convc1.ffn1.2: 42240 x 324 x 486, bias + residual + GELU + dw1), a few launches: ncu target."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
M, N, K = 42240, 324, 486
buf = lambda n, c: torch.randn((n, (c + 7) // 8 * 8), device="cuda").half()[:, :c]
x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
b = torch.randn(N, device="cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b))
torch.cuda.synchronize()
print("ok", float(o.float().abs().mean()))
