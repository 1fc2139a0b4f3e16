import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch, cases
from torch.nn.attention import sdpa_kernel, SDPBackend
from oracle import mof as omof
from video_flow_ml_b200.twins import TwinsStage01
sd = omof.mof_synthetic_weights(0, gain=1.0)
frames = [f.astype(np.float32) for f in cases.synthetic_clip(5, 288, 448, seed=5)]
for dt in (torch.float16, torch.float32):
    cnet = TwinsStage01(sd, "cnet.svt", "cuda", dt)
    for (x, y, w, h) in ((0, 0, 160, 160), (320, 160, 128, 128), (0, 0, 448, 288)):
        t = torch.stack([torch.from_numpy(f[y:y + h, x:x + w]).permute(2, 0, 1) for f in frames]).cuda()
        t = 2 * (t / 255.0) - 1.0
        ref = cnet(t)
        for b in (1, 2, 3):
            for name, be in (("default", None), ("flash", [SDPBackend.FLASH_ATTENTION]), ("efficient", [SDPBackend.EFFICIENT_ATTENTION]),
                             ("cudnn", [SDPBackend.CUDNN_ATTENTION]), ("math", [SDPBackend.MATH])):
                try:
                    if be is None:
                        o = cnet(t[:b])
                    else:
                        with sdpa_kernel(be):
                            o = cnet(t[:b])
                    print(dt, (w, h), "batch", b, name, "nan" if torch.isnan(o).any() else f"ok, vs batch-5 {float((o - ref[:b]).abs().max()):.2e}", flush=True)
                except Exception as e:
                    print(dt, (w, h), "batch", b, name, "ERR", str(e)[:80], flush=True)
