"""CPU: would storing the GMA attention matrix in fp8 hold the parity bound?

The aggregate streams one fp16 attention matrix (P x P, 99 MB at 436 x 1024) per centre frame and iteration from HBM:
297 MB of the update step's traffic.  e4m3 with one scale per row would halve that.  This script rounds the oracle's
attention matrix to fp16 / to row-scaled e4m3 and prints how far the returned flow moves (32 iterations, every weight
set of tests/golden).  Bound for a reduced-precision tier: 5e-2 px.   -> profiles/r2_fp8_attention_probe.txt
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import cases
from oracle import mof as omof
torch.set_grad_enabled(False)
orig = omof.gma_attention
def quantized(mode):
    def f(inp, sd):
        a = orig(inp, sd)
        if mode == "fp16":
            return a.half().float()
        s = 448.0 / a.amax(dim=-1, keepdim=True).clamp_min(1e-30)
        return (a * s).to(torch.float8_e4m3fn).float() / s
    return f
for gain, drift in ((1.0, None), (1.0, omof.DRIFT), (1.4, None)):
    sd = omof.mof_synthetic_weights(0, gain=gain, drift=drift) if drift else omof.mof_synthetic_weights(0, gain=gain)
    frames = cases.mof_case()
    w = omof.pad_frames(omof.frames_to_window(frames, 2, 5))
    fm, net, inp = omof.mof_encode(w, sd)
    omof.gma_attention = orig
    ref = omof.mof_refine(fm, net, inp, sd, iters=32)
    out = {}
    for mode in ("fp16", "fp8"):
        omof.gma_attention = quantized(mode)
        got = omof.mof_refine(fm, net, inp, sd, iters=32)
        out[mode] = float(torch.sqrt(((got - ref) ** 2).sum(2)).max())
    omof.gma_attention = orig
    print(f"gain {gain} drift {drift is not None}: max EPE change  fp16 attention {out['fp16']:.3e} px   fp8 e4m3 (row-scaled) attention {out['fp8']:.3e} px   |flow| max {float(ref.abs().max()):.1f}", flush=True)
