"""Prints the measured max EPE deltas of both tiers against the reference goldens (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np
import cases
from oracle import mof as omof
from video_flow_ml_b200.processing import VideoFlowProcessor

frames = cases.mof_case()
for gain, fixture in ((1.0, "mof_e2e_g10.npz"), (1.4, "mof_e2e.npz")):
    g = np.load(os.path.join(ROOT, "tests", "golden", fixture))
    for prec, lk32 in (("fp32", True), ("fp16", True), ("fp16", False)):
        p = VideoFlowProcessor("cuda", sequence_length=5, precision=prec)
        p.core.load_state_dict(omof.mof_synthetic_weights(0, gain=gain))
        p.core.model.lookup_math_f32 = lk32
        errs = []
        for idx in (2, 5):
            flow = p.compute_optical_flow(list(frames), idx)
            errs.append(float(np.sqrt(((flow - g[f"flow_{idx}"]) ** 2).sum(-1)).max()))
        print(f"weights gain {gain} tier {prec} (lookup interpolation {'fp32' if lk32 else 'fp16x2'}): max EPE delta vs reference = {max(errs):.3e} px (|flow| max {np.abs(g['flow_2']).max():.1f} px)")
        del p
