"""Generate ``tests/golden/*.npz`` by executing the UNMODIFIED reference on seeded inputs.

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py

Every array written here is an output of reference code:
  corr_lookup.npz    VideoFlow/core/Networks/MOFNetStack/corr.py  CorrBlock(...)(coords), CorrBlock.corr
  update_block.npz   MOFNetStack/stack.py SKUpdateBlock6_Deep_nopoolres_AllDecoder2.forward, gma.py Attention
  upsample.npz       MOFNetStack/network.py MOFNet.upsample_flow
  twins.npz          encoders.py twins_svt_large.forward over twins_ft.py classes (timm shim, see refshim.py)
  mof_e2e.npz        processing.VideoFlowProcessor.compute_optical_flow (full wrapper stack, 32 iterations)
  post.npz           encoding/flow_encoders.py (gamedev, rg8), effects/taa_processor.py
  host.json          prepare_frame_sequence window choice, calculate_tile_grid
Inputs are regenerated from seeds (cases.py) and weights from ``oracle.mof.mof_synthetic_weights``
(the weight generator is only a deterministic RNG; the reference consumes the result through
``load_state_dict(strict=True)``).
"""
import json
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
from oracle import mof as omof   # noqa: E402  (weight generator + manifest only)

T = torch.from_numpy


def save(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name), **arrs)
    print("wrote", name, {k: v.shape for k, v in arrs.items()})


def main():
    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    net, cfg = refshim.build_mof(iters=32)
    sd = omof.mof_synthetic_weights(0)
    net.load_state_dict(sd, strict=True)
    corr_mod = refshim.ref_module("core.Networks.MOFNetStack.corr")
    utils = refshim.ref_module("utils.utils")

    # ---- correlation volume + lookup
    f1, f2, coords = cases.lookup_case()
    blk = corr_mod.CorrBlock(T(f1), T(f2), num_levels=4, radius=4)
    out = blk(T(coords))
    vol = corr_mod.CorrBlock.corr(T(f1), T(f2))
    save("corr_lookup.npz", lookup=out.numpy(), vol0=vol[0, :2].numpy(),
         pyr3=blk.corr_pyramid[3][:64].numpy())

    # ---- one update-block call (+ attention)
    c = cases.update_case()
    Fc, _, h, w = c["net"].shape
    coords0 = utils.coords_grid(Fc, h, w)
    attn = net.att(T(c["inp"]))
    n2, m2, mask, delta = net.update_block(T(c["net"]), T(c["mhs"]), T(c["inp"]), T(c["fcorr"]), T(c["bcorr"]),
                                           T(c["fflow"]), T(c["bflow"]), coords0, attn, bs=1)
    n1, m1, mask1, delta1 = net.update_block(T(c["net"]), None, T(c["inp"]), T(c["fcorr"]), T(c["bcorr"]),
                                             T(c["fflow"]), T(c["bflow"]), coords0, attn, bs=1)
    enc = net.update_block.encoder
    pc = enc.convc1(T(c["fcorr"]))
    save("update_block.npz", attn_rows=attn[:, 0, :8].numpy(), net=n2.numpy(), mhs=m2.numpy(),
         mask=mask[:, ::37].numpy(), delta=delta.numpy(), convc1=pc.numpy(),
         net_first=n1.numpy(), mhs_first=m1.numpy(), delta_first=delta1.numpy())

    # ---- convex upsample
    fl, mk = cases.upsample_case()
    save("upsample.npz", up=net.upsample_flow(T(fl), T(mk)).numpy())

    # ---- Twins encoder (reference composition over twins_ft.py classes)
    x = cases.twins_case()
    save("twins.npz", fnet=net.fnet(T(x)).numpy()[:, ::4], cnet=net.cnet(T(x)).numpy()[:, ::4])

    # ---- end to end through the reference wrapper stack
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

            def _cfg_no_download():          # pretrain=True would hit the network
                c_ = _orig()
                c_.MOFNetStack.pretrain = False
                c_.MOFNetStack.mixed_precision = False
                return c_
            core_mod.get_cfg = _cfg_no_download
            p = proc_mod.VideoFlowProcessor("cpu", sequence_length=5)
            p.load_model()
            flows = {}
            for idx in (2, 5):
                flows[f"flow_{idx}"] = p.compute_optical_flow(list(frames), idx)
            flows["flow_u8_0"] = p.compute_optical_flow(list(cases.mof_case(as_float=False)), 0)
            win = p.prepare_frame_sequence(list(frames), 2)
            allf, _ = p.core.model(utils.InputPadder(win.shape[-2:]).pad(win), {})
            host = {"windows": {}, "tiles": {}}
            marked = [np.full((2, 2, 3), i, np.uint8) for i in range(9)]
            for n_frames in (1, 2, 4, 9):
                for Tn in (3, 5, 7, 9):
                    p.sequence_length = Tn
                    for idx in range(n_frames):
                        b = p.prepare_frame_sequence(list(marked[:n_frames]), idx)
                        host["windows"][f"{n_frames},{idx},{Tn}"] = [int(round(float(v) * 255)) for v in b[0, :, 0, 0, 0]]
            for (wd, ht) in ((1920, 1080), (3840, 2160), (1280, 1280), (1281, 100), (1024, 436), (2560, 2560)):
                r = p.calculate_tile_grid(wd, ht)
                host["tiles"][f"{wd}x{ht}"] = [r[0], r[1], r[2], r[3], r[4]]
        finally:
            os.chdir(cwd)
    save("mof_e2e.npz", all_flows_2=allf.numpy()[0, :, :, ::2, ::2], **flows)
    with open(os.path.join(HERE, "host.json"), "w") as f:
        json.dump(host, f)

    # ---- post-pass
    enc_mod = refshim.ref_module("encoding.flow_encoders")
    taa_mod = refshim.ref_module("effects.taa_processor")
    ff = cases.flow_field_case()
    hgt, wid = ff.shape[:2]
    post = {"gamedev": enc_mod.encode_flow(ff.copy(), wid, hgt, "gamedev"),
            "gamedev_4k": enc_mod.GamedevFlowEncoder().encode(ff.copy(), 3840, 2160),
            "rg8": enc_mod.MotionVectorsRG8FlowEncoder().encode(ff.copy(), wid, hgt),
            "rg8_c16": enc_mod.MotionVectorsRG8FlowEncoder(clamp_range=16.0).encode(ff.copy(), wid, hgt)}
    post["rg8_decoded"] = enc_mod.MotionVectorsRG8FlowEncoder().decode(post["rg8"])
    fr, fls = cases.taa_case()
    for name, kw in (("taa_bilateral", dict(use_flow=True, use_bilateral=True)),
                     ("taa_bilinear", dict(use_flow=True, use_bilateral=False)),
                     ("taa_simple", dict(use_flow=False, use_bilateral=False))):
        proc = taa_mod.TAAProcessor(alpha=0.1)
        prev = None
        for i, f in enumerate(fr):          # flow_processor.py:1016-1112: previous frame's flow
            o = proc.apply_taa(current_frame=f, flow_pixels=prev, previous_taa_frame=None, alpha=0.1,
                               sequence_id="s", **kw)
            post[f"{name}_{i}"] = o
            prev = fls[i].copy()
    save("post.npz", **post)


if __name__ == "__main__":
    main()
