"""Prints the MemFlow parity numbers (max EPE vs the reference golden) per step and tier.  GPU."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch
import cases
from oracle import memflow as omem, mof as omof
from video_flow_ml_b200.processing import InferenceCore, MemFlowCore
g = np.load(os.path.join(ROOT, "tests", "golden", "memflow_e2e.npz"))
frames = cases.mof_case()
x = torch.stack([torch.from_numpy(f).permute(2, 0, 1) for f in frames[:4]])[None]
x = omof.pad_frames(2 * (x / 255.0) - 1.0).contiguous().cuda()
for prec in ("fp32", "fp16"):
    core = MemFlowCore("cuda", precision=prec)
    core.load_state_dict(omem.memflow_synthetic_weights(0))
    ic = InferenceCore(core.model)
    out = []
    for t in range(3):
        low, up = ic.step(x[:, t:t + 2])
        e = np.sqrt(((up[0].cpu().numpy() - g[f"seq_up_{t}"][0]) ** 2).sum(0)).max()
        out.append(f"{e:.3e}")
    print(f"memflow tier {prec}: max EPE delta vs reference per step = {out} px (|flow| max {np.abs(g['seq_up_2']).max():.1f} px)")
