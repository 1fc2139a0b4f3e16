import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops
M, N, K = 128, 128, 64
a = torch.zeros((M, K), device="cuda", dtype=torch.float16); b = torch.zeros((N, K), device="cuda", dtype=torch.float16)
# D[m, n] = m * 128 + n  exactly in fp32 via two k columns: a[:,0]=m, b[:,0]=128 ; a[:,1]=1, b[:,1]=n
a[:, 0] = torch.arange(M, device="cuda").half(); b[:, 0] = 128
a[:, 1] = 1; b[:, 1] = torch.arange(N, device="cuda").half()
for odt in (torch.float32, torch.float16):
    out = torch.full((M, N), -1.0, device="cuda", dtype=odt)
    ops.gemm(a, b, out)
    torch.cuda.synchronize()
    want = (torch.arange(M, device="cuda").view(-1, 1) * 128 + torch.arange(N, device="cuda").view(1, -1)).float()
    if odt == torch.float16:
        want = want.half().float()
    bad = out.float() != want
    idx = bad.nonzero()
    print(odt, int(bad.sum()), "wrong; rows", sorted(set(idx[:, 0].tolist()))[:40], "cols", sorted(set(idx[:, 1].tolist()))[:80])
    for (m, n) in idx[:24].tolist():
        g = int(out[m, n]); print(f"  out[{m},{n}] = {g} = (row {g // 128}, col {g % 128})")
