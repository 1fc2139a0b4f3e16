"""One GEMM shape timed three ways (graph of 20 back to back / single eager launches / graph with an L2-sized gap), to
tell kernel time from launch-chain effects:  python scripts/micro/gemm_time_modes.py [M N K]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
dev = "cuda"
M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (42240, 486, 324)
def buf(n, c): return torch.randn((n, (c + 7) // 8 * 8), device=dev).half()[:, :c]
x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
b = torch.randn(N, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
def graph_time(fn, reps):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = ev(), ev()
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000 / reps)
    return min(ts), sorted(ts)[2]
def eager_time(fn, reps=20):
    ts = []
    for _ in range(reps):
        e0, e1 = ev(), ev()
        torch.cuda.synchronize()
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000)
    return min(ts), sorted(ts)[len(ts) // 2]
for name, fn in (("plain", lambda: ops.gemm(x, w, o)), ("bias+gelu", lambda: ops.gemm(x, w, o, bias=b, act="gelu")),
                 ("bias+res+gelu", lambda: ops.gemm(x, w, o, bias=b, res=r, act="gelu"))):
    print(f"{M}x{N}x{K} {name:14s} graph x20 (min, median) {graph_time(fn, 20)}  graph x1 {graph_time(fn, 1)}  eager single {eager_time(fn)}", flush=True)
