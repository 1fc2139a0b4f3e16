"""CPU: how far does the REFERENCE'S OWN reduced-precision path sit from its fp32 output?

The reference ships `mixed_precision=True`: encoders, attention and update block under fp16 autocast,
volume + lookup in fp32.  This script runs the pinned oracle (a) in fp32, (b) with a 1-ulp perturbation of the
encoder outputs (conditioning of the recurrence), (c) under torch's CPU autocast in fp16 and bf16 (emulation of
that shipped path, update block + attention only: same encoder outputs), on every synthetic weight set, and prints
max EPE deltas of the returned full-resolution flows.  -> profiles/r2_fp16_floor.txt
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np
import torch
import cases
from oracle import mof as omof

torch.set_grad_enabled(False)
frames = cases.mof_case()
win = omof.pad_frames(omof.frames_to_window(frames, 2, 5))
iters = int(os.environ.get("ITERS", "32"))


def epe(a, b):
    return float(torch.sqrt(((a - b) ** 2).sum(2)).max())


sets = (("g10 (gain 1.0)", dict(gain=1.0)), ("drift (gain 1.0 + drift bias)", dict(gain=1.0, drift=omof.DRIFT)),
        ("g12 + drift", dict(gain=1.2, drift=omof.DRIFT)), ("g14 (stress)", dict(gain=1.4)))
for name, kw in sets:
    sd = omof.mof_synthetic_weights(0, **kw)
    fm, net, inp = omof.mof_encode(win, sd)
    tr = []
    ref = omof.mof_refine(fm, net, inp, sd, iters=iters, trace=tr)
    lowmax = max(float(t[0].abs().max()) for t in tr[-1:]), max(float(t[1].abs().max()) for t in tr[-1:])
    ulp = omof.mof_refine(fm * (1 + 1.2e-7), net, inp, sd, iters=iters)
    f16 = omof.mof_refine(fm, net, inp, sd, iters=iters, amp=torch.float16)
    b16 = omof.mof_refine(fm, net, inp, sd, iters=iters, amp=torch.bfloat16)
    print(f"{name}: |low-res flow| max fwd {lowmax[0]:.1f} / bwd {lowmax[1]:.1f} texels, |flow| max {float(ref.abs().max()):.1f} px; "
          f"1-ulp sensitivity {epe(ulp, ref):.2e} px; reference-style fp16 autocast {epe(f16, ref):.3e} px; bf16 autocast {epe(b16, ref):.3e} px",
          flush=True)
