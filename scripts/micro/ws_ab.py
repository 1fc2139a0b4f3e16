"""Times the update block's GEMM shapes (graph of 20 launches each) under the current VFML_* environment; one line per shape:
plain / bias+gelu / bias+res+gelu+dw1 in us.   python scripts/micro/ws_ab.py [label]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
dev = "cuda"
def buf(n, c): return torch.randn((n, (c + 7) // 8 * 8), device=dev).half()[:, :c]
def gtime(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1000 / reps)
    return best
label = sys.argv[1] if len(sys.argv) > 1 else ""
shapes = ((42240, 486, 324), (42240, 324, 486), (42240, 128, 486), (21120, 768, 512), (21120, 512, 768), (21120, 600, 400), (21120, 400, 600),
          (21120, 172, 600), (21120, 384, 256), (21120, 256, 384), (21120, 192, 128), (21120, 128, 192), (21120, 128, 128))
tot = [0.0, 0.0, 0.0]
for M, N, K in shapes:
    x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
    b = torch.randn(N, device=dev)
    t = (gtime(lambda: ops.gemm(x, w, o)), gtime(lambda: ops.gemm(x, w, o, bias=b, act="gelu")),
         gtime(lambda: ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b))))
    tot = [a + c for a, c in zip(tot, t)]
    print(f"{label:14s} {M}x{N}x{K}: {t[0]:6.1f} {t[1]:6.1f} {t[2]:6.1f}", flush=True)
print(f"{label:14s} total: {tot[0]:6.1f} {tot[1]:6.1f} {tot[2]:6.1f}")
