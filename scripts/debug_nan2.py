import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch, cases
import torch.nn.functional as F
from oracle import mof as omof
from video_flow_ml_b200 import twins
sd = omof.mof_synthetic_weights(0, gain=1.0)
frames = [f.astype(np.float32) for f in cases.synthetic_clip(5, 288, 448, seed=5)]
cnet = twins.TwinsStage01(sd, "cnet.svt", "cuda", torch.float16)
t = torch.stack([torch.from_numpy(f[:160, :160]).permute(2, 0, 1) for f in frames]).cuda()
t = 2 * (t / 255.0) - 1.0
log = []
orig = {n: getattr(F, n) for n in ("conv2d", "linear", "scaled_dot_product_attention", "gelu", "layer_norm", "pad")}
orig_ln = twins.TwinsStage01._ln
def wrap(name, fn):
    def w(*a, **k):
        o = fn(*a, **k)
        log.append((name, tuple(a[0].shape), o.detach().float().clone()))
        return o
    return w
for n, fn in orig.items():
    setattr(F, n, wrap(n, fn))
def ln(self, x, name, eps):
    o = orig_ln(self, x, name, eps)
    log.append(("ln " + name, tuple(x.shape), o.detach().float().clone()))
    return o
twins.TwinsStage01._ln = ln
cnet(t[:1]); a = list(log); log.clear()
cnet(t[:2]); b = list(log); log.clear()
for (n1, s1, o1), (n2, s2, o2) in zip(a, b):
    d = float((o1 - o2[:o1.shape[0]]).abs().max()) if o1.shape[1:] == o2.shape[1:] else float("nan")
    if o1.shape[1:] != o2.shape[1:]:
        k = o1.shape[0]
        d = float((o1 - o2[:k]).abs().max())
    print(f"{n1:40s} in {str(s1):26s} batch1-vs-batch2 max diff {d:.3e}   |out| {float(o1.abs().max()):.2f}", flush=True)
