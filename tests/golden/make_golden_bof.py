"""BOF fixture from the UNMODIFIED reference BOFNet (VideoFlow/core/Networks/BOFNet/network.py), CPU fp32,
12 iterations (configs/things.py), seeded synthetic weights:   python tests/golden/make_golden_bof.py

The reference's own VideoFlowCore cannot load a BOF checkpoint (it always builds the network from
configs/multiframes_sintel_submission.py, i.e. MOFNetStack, processing/videoflow_core.py:30,101), so
the fixture calls BOFNet.forward directly on the padded 3-frame window, exactly as
compute_flow_from_tensor would (pad -> forward -> unpad -> index shape[1] // 2).

bof_e2e.npz: flows [2, 2, H, W] (2->3 then 2->1) for the windows centred on frames 2 and 4 of cases.mof_case().
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases      # noqa: E402
import refshim    # noqa: E402
from oracle import bof as obof   # noqa: E402
from oracle import mof as omof   # noqa: E402


def build_bof(iters=12):
    refshim.install()
    cfg = importlib.import_module("configs.things").get_cfg()
    cfg.BOFNet.pretrain = False
    cfg.BOFNet.mixed_precision = False
    cfg.BOFNet.decoder_depth = iters
    net = importlib.import_module("core.Networks").build_network(cfg)
    return net.eval()


def main():
    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    net = build_bof(12)
    sd = obof.bof_synthetic_weights(0)
    ref_keys = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    mine = {k: tuple(v.shape) for k, v in sd.items()}
    assert ref_keys == mine, (sorted(set(ref_keys) ^ set(mine))[:10])
    net.load_state_dict(sd, strict=True)
    utils = refshim.ref_module("utils.utils")
    frames = cases.mof_case()
    out = {}
    for idx in (2, 4):
        win = omof.frames_to_window(frames, idx, 3)
        padder = utils.InputPadder(win.shape[-2:])
        pred, _ = net(padder.pad(win), {})
        flows = padder.unpad(pred)[0]                       # [2, 2, H, W]
        out[f"flows_{idx}"] = flows.numpy()
        mine_f = omof.unpad(obof.bof_forward(omof.pad_frames(win), sd), *win.shape[-2:])[0]
        print(idx, "oracle vs reference max abs", float((mine_f - flows).abs().max()), "max |flow|", float(flows.abs().max()))
    np.savez_compressed(os.path.join(HERE, "bof_e2e.npz"), **out)
    print("wrote bof_e2e.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
