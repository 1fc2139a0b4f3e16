"""fp16 tier vs fp32 tier stage by stage on a window with repeated frames (clip start: [0, 0, 0, 1, 2])."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch, cases
from oracle import mof as omof
from video_flow_ml_b200 import ops
from video_flow_ml_b200.mof import MOFEngine
torch.set_grad_enabled(False)
sd = omof.mof_synthetic_weights(0, gain=1.0)
frames = [f.astype(np.float32)[:160, :160] for f in cases.synthetic_clip(5, 288, 448, seed=5)]
for idx in (0, 2):
    win = omof.pad_frames(omof.frames_to_window(frames, idx, 5)).cuda()
    e32 = MOFEngine(sd, "cuda", torch.float32, iters=4, use_graph=False)
    e16 = MOFEngine(sd, "cuda", torch.float16, iters=4, encoder_dtype=torch.float16, use_graph=False)
    fm, net, inp = e32.encode(win)
    fm16, net16, inp16 = e16.encode(win)
    print(f"idx {idx}: fmaps fp16-enc vs fp32-enc max {float((fm16.float() - fm).abs().max()):.3e} (|f| max {float(fm.abs().max()):.2f})")
    perm = ops.corr_lookup_perm("cuda")
    res = {}
    for name, eng, f_ in (("fp32", e32, fm), ("fp16", e16, fm)):          # SAME fp32-encoder features into both tiers
        T, _, h, w = f_.shape
        Fc, P = T - 2, h * w
        ws = eng._workspace(Fc, h, w)
        fpyr, bpyr = eng.build_pyramids(f_.float().contiguous(), ws)
        ops.nchw_to_rows(net.float().contiguous(), ws["x512"][:, 0:128]); ops.nchw_to_rows(inp.float().contiguous(), ws["x512"][:, 128:256])
        ops.rows_fill(ws["x400"][:, 352:400], eng.init_mhs)
        eng.attention(ws, Fc, P)
        fc = ops.coords_grid(Fc, h, w, eng.dev); bc = ops.coords_grid(Fc, h, w, eng.dev)
        trace = []
        for it in range(4):
            eng.update_step(ws, fpyr, bpyr, fc, bc, Fc, h, w)
            c2 = ws["corr2"].float().clone()
            if eng.tiled:
                c2r = torch.empty_like(c2); c2r[:, perm] = c2; c2 = c2r
            trace.append((c2, ws["cor_in"].float().clone(), ws["x512"].float().clone(), fc.clone(), bc.clone(), ws["delta"].float().clone()))
        res[name] = trace
        attn = ws["attn"].float().clone() if name == "fp16" else None
    for it in range(4):
        a, b = res["fp32"][it], res["fp16"][it]
        names = ("lookup", "cor_in", "x512", "fcoords", "bcoords", "delta")
        print(f"  iter {it}: " + "  ".join(f"{n} {float((x - y).abs().max()):.2e}/{float(x.abs().max()):.1f}" for n, x, y in zip(names, a, b)))
    d = (res["fp32"][0][0] - res["fp16"][0][0]).abs()
    r, c = divmod(int(d.argmax()), d.shape[1])
    print(f"  worst lookup element iter 0: row {r} (dir {r // (3 * 400)}, frame {(r // 400) % 3}, pixel {r % 400}), channel {c} (level {c // 81}, i {c % 81 // 9}, j {c % 9}): fp32 {float(res['fp32'][0][0][r, c]):.4f} fp16 {float(res['fp16'][0][0][r, c]):.4f}")
