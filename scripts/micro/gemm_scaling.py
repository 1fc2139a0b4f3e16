"""Per-CTA throughput of the tcgen05 GEMM against the number of CTAs (VFML_GEMM_MAX_CTAS): shared-resource (L2) contention
or per-SM latency?  Run once per setting: the cap is read at first use."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from video_flow_ml_b200 import ops
dev = "cuda"
PITCH = int(os.environ.get("VFML_PITCH", "8"))     # row pitch granularity in elements (16 -> 32-byte aligned rows)
def buf(n, c): return torch.randn((n, (c + PITCH - 1) // PITCH * PITCH), device=dev).half()[:, :c]
def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1000.0
cap = os.environ.get("VFML_GEMM_MAX_CTAS", "148")
out = []
for M, N, K in ((42240, 486, 324), (21120, 768, 512), (21120, 512, 768), (42240, 128, 486)):
    x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
    b = torch.randn(N, device=dev)
    t_plain = timeit(lambda: ops.gemm(x, w, o, weights=True))
    t_gelu = timeit(lambda: ops.gemm(x, w, o, bias=b, act="gelu", weights=True))
    t_res = timeit(lambda: ops.gemm(x, w, o, bias=b, res=r, act="gelu", weights=True))
    out.append(f"{M}x{N}x{K}: plain {t_plain:6.1f} gelu {t_gelu:6.1f} res+gelu {t_res:6.1f} us")
print(f"CTAs <= {cap:>3s} pitch%{PITCH} | " + " | ".join(out))
