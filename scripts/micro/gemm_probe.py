"""Where does the tensor-core GEMM's time go?  Same shape, different epilogues / K.  Run on the GPU box."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops

dev = "cuda"
def buf(n, c, dtype=torch.float16):
    return torch.randn((n, (c + 7) // 8 * 8), device=dev).to(dtype)[:, :c]

def timeit(fn, reps=50):
    for _ in range(5): fn()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1000.0

for M, N, K in ((42240, 486, 324), (42240, 512, 324), (42240, 256, 324), (42240, 486, 64), (21120, 768, 512), (21120, 192, 128), (21120, 128, 128)):
    x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
    b = torch.randn(N, device=dev)
    row = [f"{M}x{N}x{K}"]
    for label, fn in (("plain", lambda: ops.gemm(x, w, o)), ("bias", lambda: ops.gemm(x, w, o, bias=b)),
                      ("bias+gelu", lambda: ops.gemm(x, w, o, bias=b, act="gelu")),
                      ("bias+res+gelu", lambda: ops.gemm(x, w, o, bias=b, res=r, act="gelu")),
                      ("bias+res+gelu+dw1", lambda: ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b)))):
        row.append(f"{label} {timeit(fn):6.1f}")
    print("  ".join(row))
