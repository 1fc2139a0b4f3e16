"""Per-op timing of one MOF update iteration at config C1 (F=3, 55x128), fp16 tier, CUDA events, warm.
Run on the GPU box:  python scripts/microbench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops

dev = "cuda"
Fc, h, w = 3, 55, 128
P = h * w
Np = Fc * P
dt = torch.float16


def buf(n, c, dtype=dt):
    return torch.randn((n, (c + 7) // 8 * 8), device=dev).to(dtype)[:, :c]


def timeit(fn, reps=20):
    """us per call, replayed from a CUDA graph (an eager ctypes launch costs ~12 us of CPU time, more than the small kernels)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1000.0   # us


total = 0.0
print(f"{'op':44s} {'us':>9s} {'TFLOP/s':>9s}")
# PCBlock GEMMs: (name, M, Cin, Cout)
blocks = [("convc1", 2 * Np, 324, 128, 15), ("convc2", Np, 256, 192, 15), ("convf2", Np, 128, 64, 15),
          ("conv", Np, 400, 172, 15), ("gru", Np, 512, 128, 7), ("fhead", Np, 128, 4, 15)]
for name, M, C, Co, k in blocks:
    hid = int(1.5 * C)
    x, t1, x2, x3, out = buf(M, C), buf(M, hid), buf(M, C), buf(M, C), buf(M, Co)
    w1a, w1b, wpw, w2a, w2b = buf(hid, C), buf(C, hid), buf(C, C), buf(hid, C), buf(Co, hid)
    bh, bc, bo = torch.randn(hid, device=dev), torch.randn(C, device=dev), torch.randn(Co, device=dev)
    dww = torch.randn((k * k, C), device=dev) / k
    dwp = ops.dwconv_pack_weights(dww, k)
    ops_list = [
        (f"{name}.ffn1.0 {M}x{hid}x{C} gelu", lambda: ops.gemm(x, w1a, t1, bias=bh, act="gelu"), 2.0 * M * hid * C),
        (f"{name}.ffn1.2 {M}x{C}x{hid} res gelu dw1", lambda: ops.gemm(t1, w1b, x2, bias=bc, res=x, act="gelu", dw=(bc, bc)), 2.0 * M * hid * C),
        (f"{name}.dw{k} C={C} (tensor cores)", lambda: ops.dwconv_tc_res_gelu(x2, x3, M // P, h, w, dwp, bc, k), 2.0 * M * C * k * k),
        (f"{name}.pw {M}x{C}x{C} res gelu", lambda: ops.gemm(x3, wpw, x2, bias=bc, res=x3, act="gelu"), 2.0 * M * C * C),
        (f"{name}.ffn2.0 {M}x{hid}x{C} gelu", lambda: ops.gemm(x2, w2a, t1, bias=bh, act="gelu"), 2.0 * M * hid * C),
        (f"{name}.ffn2.2 {M}x{Co}x{hid}", lambda: ops.gemm(t1, w2b, out, bias=bo), 2.0 * M * hid * Co),
    ]
    for label, fn, flops in ops_list:
        us = timeit(fn)
        total += us
        print(f"{label:44s} {us:9.1f} {flops / us / 1e6:9.1f}")
# aggregate: one affinity matrix per centre frame (3 x 99 MB: cycles through HBM, as in the engine)
attns = [torch.softmax(torch.randn((P, P), device=dev), -1).to(dt) for _ in range(Fc)]
vT, mf, o = buf(128, P), buf(P, 128), buf(P, 128)
def agg_single():
    for a_ in attns:
        ops.gemm(a_, vT, o, res=mf, alpha=0.5)
us = timeit(agg_single) / Fc
print(f"{'aggregate attn.v (per frame) 7040x128x7040':44s} {us:9.1f} {2.0 * P * P * 128 / us / 1e6:9.1f}   x{Fc}  single pass")
for S in (2, 5, 10):
    part = torch.empty(S * P * 128, device=dev, dtype=torch.float32)
    def agg_split():
        for a_ in attns:
            ops.gemm_splitk(a_, vT, o, S, part, res=mf, alpha=0.5)
    us_s = timeit(agg_split) / Fc
    print(f"{'  split-K ' + str(S) + ' + reduce':44s} {us_s:9.1f} {2.0 * P * P * 128 / us_s / 1e6:9.1f}   ({P * P * 2 / us_s / 1e6:.2f} TB/s of affinity rows)")
    if S == 5:
        us = us_s
total += Fc * us
wv = buf(128, 128)
us = timeit(lambda: ops.gemm(wv, mf, vT))
print(f"{'to_v^T (per frame) 128x7040x128':44s} {us:9.1f}   x{Fc}")
total += Fc * us
# lookup pair + glue
f = torch.randn((Fc, 256, h, w), device=dev)
fp = ops.corr_volume_pyramid_tiled(f, torch.randn_like(f))
bp = ops.corr_volume_pyramid_tiled(f, torch.randn_like(f))
us = timeit(lambda: ops.corr_volume_pyramid_tiled(f, f), 5)
print(f"{'corr_volume_pyramid (3 pairs)':44s} {us:9.1f} {2.0 * Fc * P * P * 256 / us / 1e6:9.1f}   (once per direction per window)")
coords = ops.coords_grid(Fc, h, w, dev) + 4 * torch.randn((Fc, 2, h, w), device=dev)
corr2 = buf(2 * Np, 324)
us = timeit(lambda: ops.corr_lookup_tiled(fp, bp, coords, coords, corr2[:Np], corr2[Np:]))
print(f"{'corr_lookup_tiled (same coords: L2-warm)':44s} {us:9.1f}")
total += us
x400, f128, mfb = buf(Np, 400), buf(Np, 128), buf(Np, 128)
w1, b1 = torch.randn((128, 4), device=dev), torch.randn(128, device=dev)
us = timeit(lambda: ops.motion_glue(coords, coords, x400, f128, mfb, w1, b1, Fc, h, w))
print(f"{'motion_glue':44s} {us:9.1f}")
total += us
print(f"sum of one iteration's ops: {total:.1f} us  -> x32 = {total * 32 / 1000:.1f} ms")
