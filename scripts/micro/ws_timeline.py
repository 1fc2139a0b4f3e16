import sys, torch
sys.path.insert(0, ".")
from video_flow_ml_b200 import ops
M, N, K = (int(a) for a in sys.argv[1:4])
buf = lambda n, c: torch.randn((n, (c + 7) // 8 * 8), device="cuda").half()[:, :c]
x, w, o = buf(M, K), buf(N, K), buf(M, N)
for _ in range(3): ops.gemm(x, w, o)
torch.cuda.synchronize()
