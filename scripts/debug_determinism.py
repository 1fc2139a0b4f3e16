"""Debug helper: are two evaluations of the same tile bit-identical?  (run on the GPU box)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch
import cases
from oracle import mof as omof
from video_flow_ml_b200.processing import VideoFlowProcessor

p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp32")
p.core.load_state_dict(omof.mof_synthetic_weights(0), iters=4)
frames = cases.synthetic_clip(5, 192, 352, seed=5)
tile = [f[0:96, 0:96] for f in frames]
eng = p.core.model
w1 = p.prepare_frame_sequence(tile, 2)
w2 = p.prepare_frame_sequence(frames, 2)[..., 0:96, 0:96].contiguous()
print("window equal:", torch.equal(w1, w2))
a = eng.encode(w1); b = eng.encode(w2); c = eng.encode(w1)
for n, x, y, z in zip(("fmaps", "net", "inp"), a, b, c):
    print(n, "path diff", (x - y).abs().max().item(), "repeat diff", (x - z).abs().max().item())
r1 = eng.refine(*a); r2 = eng.refine(*a)
print("refine repeat diff", (r1 - r2).abs().max().item())
other = eng.refine(*eng.encode(p.prepare_frame_sequence([f[0:96, 288:352] for f in frames], 2)))
r3 = eng.refine(*a)
print("refine after other-shape diff", (r1 - r3).abs().max().item())
fp1, bp1 = eng.build_pyramids(a[0]); fp2, bp2 = eng.build_pyramids(a[0])
print("pyramid repeat diff", max((x - y).abs().max().item() for x, y in zip(fp1 + bp1, fp2 + bp2)))
