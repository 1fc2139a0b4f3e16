"""Round-2 fixtures from the UNMODIFIED reference (CPU fp32):   python tests/golden/make_golden_r2.py [names...]

  mof_e2e_drift.npz  large-motion fixture: gain-1.0 weights + a drift in update_block.flow_head.ffn2.2.bias
                     (oracle.mof.DRIFT), 124x188, 32 iterations through processing.VideoFlowProcessor: the low-res
                     coordinates travel ~20 texels (160 px) and leave the image, so the recurrent lookups run far
                     from the identity grid (in range, border, out of range).  Also stores the conditioning of the
                     fixture measured on the pinned oracle: `ulp_sens` (output change for a 1-ulp perturbation of
                     the encoder outputs) and `floor_fp16` (deviation of the reference-style fp16 autocast path,
                     oracle.mof._amp, from the fp32 output) -- the noise floor of the reference's OWN shipped
                     mixed-precision path on this fixture.
  mof_c1.npz         BASELINE config C1 size: 436x1024, T=5, 4 iterations (drift weights), wrapper stack.
  mof_c3.npz         BASELINE config C3 size: 376x1242, T=9, 4 iterations, wrapper stack.
  bof_tile.npz       BASELINE config C2 tile: BOFNet on one 1280x1080 tile (P = 21600), 3 frames, 4 iterations.
  mof_tiled.npz      processing.VideoFlowProcessor.compute_optical_flow_tiled (reference :231-283) with
                     calculate_tile_grid's tile_size default set to 160 on a 288x448 clip (2x3 tiles, ragged: 160/128 wide and high;
                     every tile keeps >= 16 feature texels per side, below that the reference's own level-3 grid_sample
                     divides by W-1 = 0 and returns NaN), 4 iterations.

Full-size outputs are stored subsampled with stride 3 (coprime with the 8x8 upsampling cell, so every one of the 64
sub-pixel mask positions is covered).  Inputs are regenerated from seeds (cases.py), weights from the oracle's
deterministic generator; the reference consumes them through its own load path (torch.load + load_state_dict).
"""
import contextlib
import importlib
import os
import sys
import tempfile
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases      # noqa: E402
import refshim    # noqa: E402
from oracle import bof as obof   # noqa: E402
from oracle import mof as omof   # noqa: E402

SUB = 3


def save(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name), **arrs)
    print("wrote", name, {k: getattr(v, "shape", v) for k, v in arrs.items()}, flush=True)


@contextlib.contextmanager
def ref_processor(sd, iters, T):
    """The reference's VideoFlowProcessor('cpu') loaded with `sd` through its own load_model()."""
    refshim.build_mof(iters=iters)
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.symlink(os.path.join(refshim.REF, "VideoFlow"), os.path.join(tmp, "VideoFlow"))
        os.makedirs(os.path.join(tmp, "VideoFlow_ckpt"))
        torch.save({"module." + k: v for k, v in sd.items()}, os.path.join(tmp, "VideoFlow_ckpt", "MOF_sintel.pth"))
        os.chdir(tmp)
        try:
            proc_mod = refshim.ref_module("processing.videoflow_processor")
            core_mod = refshim.ref_module("processing.videoflow_core")
            orig = getattr(core_mod, "_vfml_orig_get_cfg", None) or core_mod.get_cfg
            core_mod._vfml_orig_get_cfg = orig

            def cfg_no_download():
                c_ = orig()
                c_.MOFNetStack.pretrain = False
                c_.MOFNetStack.mixed_precision = False
                c_.MOFNetStack.decoder_depth = iters
                return c_
            core_mod.get_cfg = cfg_no_download
            p = proc_mod.VideoFlowProcessor("cpu", sequence_length=T)
            p.load_model()
            yield p
        finally:
            os.chdir(cwd)


def drift():
    sd = omof.mof_synthetic_weights(0, gain=1.0, drift=omof.DRIFT)
    frames = cases.mof_case()
    utils = refshim.ref_module("utils.utils") if False else None
    with ref_processor(sd, 32, 5) as p:
        flows = {f"flow_{i}": p.compute_optical_flow(list(frames), i) for i in (2, 5)}
        win = p.prepare_frame_sequence(list(frames), 2)
        utils = refshim.ref_module("utils.utils")
        allf, _ = p.core.model(utils.InputPadder(win.shape[-2:]).pad(win), {})
    # conditioning of the fixture, on the oracle (pinned to the reference by test_oracle_golden.py)
    w = omof.pad_frames(omof.frames_to_window(frames, 2, 5))
    fm, net, inp = omof.mof_encode(w, sd)
    ref = omof.mof_refine(fm, net, inp, sd, iters=32)
    e = lambda a: float(torch.sqrt(((a - ref) ** 2).sum(2)).max())
    ulp = e(omof.mof_refine(fm * (1 + 1.2e-7), net, inp, sd, iters=32))
    f16 = e(omof.mof_refine(fm, net, inp, sd, iters=32, amp=torch.float16))
    print("oracle vs reference (all flows):", float((ref - allf).abs().max()), "ulp", ulp, "fp16 floor", f16)
    save("mof_e2e_drift.npz", all_flows_2=allf.numpy()[0, :, :, ::2, ::2], ulp_sens=np.float64(ulp),
         floor_fp16=np.float64(f16), **flows)


def c1():
    sd = omof.mof_synthetic_weights(0, gain=1.0, drift=omof.DRIFT)
    frames = [f.astype(np.float32) for f in cases.synthetic_clip(8, 436, 1024, seed=1234)]
    t0 = time.time()
    with ref_processor(sd, 4, 5) as p:
        flow = p.compute_optical_flow(list(frames), 3)
    print("C1 reference window (4 iterations):", time.time() - t0, "s; max |flow|", float(np.abs(flow).max()))
    save("mof_c1.npz", flow_3=flow[::SUB, ::SUB].copy())


def c3():
    sd = omof.mof_synthetic_weights(0, gain=1.0, drift=omof.DRIFT)
    frames = [f.astype(np.float32) for f in cases.synthetic_clip(10, 376, 1242, seed=4321)]
    t0 = time.time()
    with ref_processor(sd, 4, 9) as p:
        flow = p.compute_optical_flow(list(frames), 4)
    print("C3 reference window (4 iterations):", time.time() - t0, "s; max |flow|", float(np.abs(flow).max()))
    save("mof_c3.npz", flow_4=flow[::SUB, ::SUB].copy())


def bof_tile():
    refshim.install()
    cfg = importlib.import_module("configs.things").get_cfg()
    cfg.BOFNet.pretrain = False
    cfg.BOFNet.mixed_precision = False
    cfg.BOFNet.decoder_depth = 4
    net = importlib.import_module("core.Networks").build_network(cfg).eval()
    sd = obof.bof_synthetic_weights(0)
    net.load_state_dict(sd, strict=True)
    utils = refshim.ref_module("utils.utils")
    # the left tile (1280 x 1080) of a 1920x1080 clip (calculate_tile_grid: tiles 1280x1080 + 640x1080)
    frames = [f.astype(np.float32)[:, :1280] for f in cases.synthetic_clip(3, 1080, 1920, seed=99)]
    win = omof.frames_to_window(frames, 1, 3)
    padder = utils.InputPadder(win.shape[-2:])
    t0 = time.time()
    pred, _ = net(padder.pad(win), {})
    flows = padder.unpad(pred)[0].numpy()            # [2, 2, H, W]: 2->3, 2->1
    print("BOF 1280x1080 tile (4 iterations):", time.time() - t0, "s; max |flow|", float(np.abs(flows).max()))
    save("bof_tile.npz", flows_1=flows[:, :, ::5, ::5].copy())        # stride 5: also coprime with 8


def tiled():
    sd = omof.mof_synthetic_weights(0, gain=1.0)
    frames = cases.synthetic_clip(5, 288, 448, seed=5)           # uint8 (the /255 branch)
    ffl = [f.astype(np.float32) for f in frames]                  # float frames: input-driven features
    with ref_processor(sd, 4, 5) as p:
        orig = p.calculate_tile_grid
        p.calculate_tile_grid = lambda w, h, tile_size=160: orig(w, h, tile_size)
        p.set_tile_mode(True)
        out = {"tiled_u8_2": p.compute_optical_flow_tiled(list(frames), 2)[::SUB, ::SUB].copy(),
               "tiled_f32_2": p.compute_optical_flow_tiled(list(ffl), 2),
               "tiled_f32_0": p.compute_optical_flow_tiled(list(ffl), 0)[::SUB, ::SUB].copy()}
    save("mof_tiled.npz", **out)


ALL = {"drift": drift, "c1": c1, "c3": c3, "bof_tile": bof_tile, "tiled": tiled}

if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    for n in (sys.argv[1:] or list(ALL)):
        ALL[n]()
