"""MemFlow fixture from the UNMODIFIED reference MemFlowNet + InferenceCore (MemFlow/core/Networks/MemFlowNet/MemFlow.py,
MemFlow/inference/inference_core_skflow.py), CPU fp32, 15 iterations (configs/sintel_memflownet.py), seeded synthetic
weights:   python tests/golden/make_golden_memflow.py

Runs in its own process: MemFlow's top-level packages (`core`, `configs`, `inference`, `utils`) shadow VideoFlow's.
The flash-attention branch of MemoryManager.match_memory (bf16, CUDA only) is disabled by setting the module flag
FLASH_AVAIABLE = False after import, which selects the reference's own einsum + do_softmax path
(memory_manager_skflow.py:58-61,88-89).

memflow_e2e.npz:
  seq_low_{t}, seq_up_{t}   InferenceCore.step outputs for t = 0, 1, 2 of one clip (memory empty / 1 frame / replaced)
  core_flow                 what MemFlowCore.compute_flow_from_tensor returns for the 3-frame window centred on frame 2
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
from oracle import memflow as omem   # noqa: E402
from oracle import mof as omof       # noqa: E402

REF = os.path.join(refshim.REF, "MemFlow")


def build_memflow():
    refshim._mod("yacs")
    refshim._mod("yacs.config", CfgNode=refshim.CfgNode)
    timm = refshim._mod("timm", _vfml_shim=True)      # encoders.py imports timm for the (unused) twins option
    timm.__path__ = []
    for p in (os.path.join(REF, "inference"), os.path.join(REF, "core"), REF):
        sys.path.insert(0, p)
    cfg = importlib.import_module("configs.sintel_memflownet").get_cfg()
    net = importlib.import_module("core.Networks").build_network(cfg).eval()
    core_mod = importlib.import_module("inference.inference_core_skflow")
    mm = importlib.import_module("inference.memory_manager_skflow")
    mm.FLASH_AVAIABLE = False
    return net, cfg, core_mod.InferenceCore


def clip_tensor(frames):
    """float frames holding 0..255 -> [1, T, 3, H, W] in [-1, 1] (memflow_inference_isolated.py:64-66)"""
    x = torch.stack([torch.from_numpy(f).permute(2, 0, 1) for f in frames])[None]
    return 2 * (x / 255.0) - 1.0


def main():
    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    net, cfg, InferenceCore = build_memflow()
    sd = omem.memflow_synthetic_weights(0)
    ref_keys = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    mine = {k: tuple(v.shape) for k, v in sd.items()}
    assert ref_keys == mine, sorted(set(ref_keys.items()) ^ set(mine.items()))[:10]
    net.load_state_dict(sd, strict=True)
    utils = importlib.import_module("utils.utils")
    frames = cases.mof_case()
    out = {}

    # (1) a sequence through one InferenceCore: the memory is empty, then holds one frame, then is replaced
    x = clip_tensor(frames[:4])
    padder = utils.InputPadder(x.shape)
    xp = padder.pad(x)
    ref = InferenceCore(net, config=cfg)
    mine_core = omem.InferenceCore(sd)
    for t in range(3):
        low, up = ref.step(xp[:, t:t + 2], end=False, add_pe=False, flow_init=None)
        out[f"seq_low_{t}"], out[f"seq_up_{t}"] = low.numpy(), up.numpy()
        mlow, mup = mine_core.step(xp[:, t:t + 2])
        print(t, "oracle vs reference: low", float((mlow - low).abs().max()), "up", float((mup - up).abs().max()),
              "max |flow|", float(up.abs().max()), "memory", ref.memory.work_mem.size)

    # (2) the application's call: fresh core, last two frames of the window, end=True
    win = omof.frames_to_window(frames, 2, 3)            # float frames pass through unscaled: values 0..255
    xs = 2 * (win / 255.0) - 1.0
    padder = utils.InputPadder(xs.shape)
    core = InferenceCore(net, config=cfg)
    _, pred = core.step(padder.pad(xs)[:, -2:], end=True, add_pe=False, flow_init=None)
    flow = padder.unpad(pred[0])
    out["core_flow"] = flow.numpy()
    mine_flow = omem.core_compute_flow(win, sd)
    print("core: oracle vs reference", float((mine_flow - flow).abs().max()), "max |flow|", float(flow.abs().max()))
    np.savez_compressed(os.path.join(HERE, "memflow_e2e.npz"), **out)
    print("wrote memflow_e2e.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
