"""Second end-to-end fixture from the UNMODIFIED reference wrapper stack, with the well-conditioned
synthetic weight set (update-block gain 1.0):   python tests/golden/make_golden_wellcond.py

Why a second set: with the gain-1.4 weights of mof_e2e.npz the 32-iteration recurrence amplifies a
1-ulp (1e-7 relative) perturbation of the encoder outputs to 3e-3 px in the reference's own CPU
output, so a 1e-2 px bound sits at the fp32 noise floor of ANY re-ordered fp32 implementation.
With gain 1.0 the same perturbation moves the output by 3e-5 px, which makes <= 1e-2 px a
meaningful bound for the fp32 tier and <= 5e-2 px one for the fp16 tier.

mof_e2e_g10.npz: processing.VideoFlowProcessor.compute_optical_flow (idx 2, 5) and the full model
output for the window at idx 2, reference code, CPU fp32, 32 iterations.
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases      # noqa: E402
import refshim    # noqa: E402
from oracle import mof as omof   # noqa: E402  (weight generator only)

GAIN = 1.0


def main():
    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    refshim.build_mof(iters=32)          # installs the import shims
    sd = omof.mof_synthetic_weights(0, gain=GAIN)
    utils = refshim.ref_module("utils.utils")
    frames = cases.mof_case()
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.symlink(os.path.join(refshim.REF, "VideoFlow"), os.path.join(tmp, "VideoFlow"))
        os.makedirs(os.path.join(tmp, "VideoFlow_ckpt"))
        torch.save({"module." + k: v for k, v in sd.items()}, os.path.join(tmp, "VideoFlow_ckpt", "MOF_sintel.pth"))
        os.chdir(tmp)
        try:
            proc_mod = refshim.ref_module("processing.videoflow_processor")
            core_mod = refshim.ref_module("processing.videoflow_core")
            _orig = core_mod.get_cfg

            def _cfg_no_download():
                c_ = _orig()
                c_.MOFNetStack.pretrain = False
                c_.MOFNetStack.mixed_precision = False
                return c_
            core_mod.get_cfg = _cfg_no_download
            p = proc_mod.VideoFlowProcessor("cpu", sequence_length=5)
            p.load_model()
            flows = {f"flow_{i}": p.compute_optical_flow(list(frames), i) for i in (2, 5)}
            win = p.prepare_frame_sequence(list(frames), 2)
            allf, _ = p.core.model(utils.InputPadder(win.shape[-2:]).pad(win), {})
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(HERE, "mof_e2e_g10.npz"), all_flows_2=allf.numpy()[0, :, :, ::2, ::2], **flows)
    print("wrote mof_e2e_g10.npz", {k: v.shape for k, v in flows.items()}, "max |flow|", float(np.abs(allf.numpy()).max()))


if __name__ == "__main__":
    main()
