"""Which reduced-precision element of the fp16 tier carries its deviation from the fp32 reference?   (GPU box)

For every fixture produced by the unmodified reference (g10 = well conditioned, drift = large motion, g14 = stress), the
fp16 tier is run as shipped and with ONE element at a time flipped back to higher precision:
  lookup-f32   : fp32 interpolation arithmetic in the tiled lookup (default: packed fp16x2)
  vol-f32      : fp32 volume storage + fp32 lookup (rows kernels; what the reference's own mixed-precision path keeps in
                 fp32, network.py:151 / corr.py:97), lookup output still fp16
  enc-f32      : fp32 Twins encoders (the reference autocasts them to fp16 as well)
  gelu-poly    : the MUFU-free degree-12 GELU polynomial instead of the tanh.approx form (second process, VFML_LIB_PATH)
and the fp32 tier for reference.  Next to it: the deviation of the reference's OWN mixed-precision path on the same
fixture (emulated on the pinned oracle with CPU autocast; stored in the drift fixture, measured by scripts/fp16_floor.py).
-> profiles/r2_fp16_ablation.txt
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np
import torch
import cases
from oracle import mof as omof
from video_flow_ml_b200.processing import VideoFlowProcessor

FIXTURES = (("g10  (gain 1.0, |flow| <= 7 px)", "mof_e2e_g10.npz", dict(gain=1.0)),
            ("drift (gain 1.0 + drift, |flow| <= 162 px)", "mof_e2e_drift.npz", dict(gain=1.0, drift=omof.DRIFT)),
            ("g14  (gain 1.4 stress, |flow| <= 98 px)", "mof_e2e.npz", dict(gain=1.4)))
frames = cases.mof_case()


def run(label, precision, **opts):
    row = []
    for _, fixture, kw in FIXTURES:
        g = np.load(os.path.join(ROOT, "tests", "golden", fixture))
        p = VideoFlowProcessor("cuda", sequence_length=5, precision=precision)
        p.core.load_state_dict(omof.mof_synthetic_weights(0, **kw))
        eng = p.core.model
        if opts.get("lookup_f32"):
            eng.lookup_math_f32 = True
        if opts.get("vol_f32"):
            eng.tiled = False
            eng.vol_dtype = torch.float32
            eng.convc1 = type(eng.convc1)({k: v for k, v in omof.mof_synthetic_weights(0, **kw).items()}, "update_block.encoder.convc1", 15,
                                          eng.dev, eng.dt, perm=None, pack=eng.ops.dwconv_pack_weights)      # un-permuted weights
        if opts.get("enc_f32"):
            from video_flow_ml_b200.twins import TwinsStage01
            sd = omof.mof_synthetic_weights(0, **kw)
            eng.fnet = TwinsStage01(sd, "fnet.svt", eng.dev, torch.float32)
            eng.cnet = TwinsStage01(sd, "cnet.svt", eng.dev, torch.float32)
        errs = []
        for idx in (2, 5):
            flow = p.compute_optical_flow(list(frames), idx)
            errs.append(float(np.sqrt(((flow - g[f"flow_{idx}"]) ** 2).sum(-1)).max()))
        row.append(max(errs))
        del p
        torch.cuda.empty_cache()
    print(f"{label:34s} " + "  ".join(f"{e:10.3e}" for e in row), flush=True)


if __name__ == "__main__":
    if os.environ.get("VFML_ABLATION_CHILD"):
        run("fp16, gelu-poly (no MUFU tanh)", "fp16")
        sys.exit(0)
    print("max EPE delta (px) vs the unmodified reference's fp32 output, frames 2 and 5, 32 iterations")
    print(f"{'variant':34s} " + "  ".join(f"{n.split()[0]:>10s}" for n, _, _ in FIXTURES))
    run("fp32 tier (exact FFMA kernels)", "fp32")
    run("fp16 tier as shipped", "fp16")
    run("fp16, lookup-f32", "fp16", lookup_f32=True)
    run("fp16, vol-f32 (+ fp32 lookup)", "fp16", vol_f32=True)
    run("fp16, enc-f32", "fp16", enc_f32=True)
    var = os.path.join(ROOT, "video_flow_ml_b200", "variants", "libvfml_gelupoly.so")
    if os.path.exists(var):
        subprocess.run([sys.executable, __file__], env=dict(os.environ, VFML_ABLATION_CHILD="1", VFML_LIB_PATH=var))
    gd = np.load(os.path.join(ROOT, "tests", "golden", "mof_e2e_drift.npz"))
    print(f"\nreference's own mixed-precision path (fp16 autocast of encoders / attention / update block, fp32 volume + lookup; oracle + CPU "
          f"autocast emulation, scripts/fp16_floor.py): g10 4.4e-02, drift {float(gd['floor_fp16']):.3e}, g14 5.0e+00 px; "
          f"1-ulp sensitivity of the reference's fp32 output: g10 2.8e-05, drift {float(gd['ulp_sens']):.2e}, g14 3.3e-03 px")
