"""Correlation-lookup roofline at config C1 (F=3, 55x128, both directions = one launch), fp16 tier.
Rotates 8 coordinate sets so successive launches touch different texels of the ~0.9 GB of volumes.
Run on the GPU box:  python scripts/lookup_bench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from video_flow_ml_b200 import ops

from video_flow_ml_b200 import _lib
dev = "cuda"
torch.zeros(1, device=dev)
Fc, h, w = 3, 55, 128
P = h * w
Np = Fc * P
g = torch.Generator(device="cpu").manual_seed(7)
f = [torch.randn((Fc, 256, h, w), generator=g).to(dev) for _ in range(3)]
tp_f = ops.corr_volume_pyramid_tiled(f[0], f[1])
tp_b = ops.corr_volume_pyramid_tiled(f[0], f[2])
py_f = ops.corr_volume_pyramid(f[0], f[1], 4, torch.float16)
py_b = ops.corr_volume_pyramid(f[0], f[2], 4, torch.float16)
lo = torch.tensor([4.0, 4.0]).view(1, 2, 1, 1)
span = torch.tensor([w - 9.0, h - 9.0]).view(1, 2, 1, 1)
rnd = [((lo + span * torch.rand((Fc, 2, h, w), generator=g)).to(dev), (lo + span * torch.rand((Fc, 2, h, w), generator=g)).to(dev)) for _ in range(8)]
grid = ops.coords_grid(Fc, h, w, dev)
smooth = [((grid + 6 * torch.randn((Fc, 2, 1, 1), generator=g).to(dev) + 0.5 * torch.randn((Fc, 2, h, w), generator=g).to(dev)).contiguous(),
           (grid + 6 * torch.randn((Fc, 2, 1, 1), generator=g).to(dev) + 0.5 * torch.randn((Fc, 2, h, w), generator=g).to(dev)).contiguous()) for _ in range(8)]
out = torch.zeros((2 * Np, 328), dtype=torch.float16, device=dev)[:, :324]
alg = 2 * Np * (4 * 100 * 2 + 8 + 324 * 2)


def timeit(fn, sets, reps=48, graph=True):
    """us per launch; the launches are replayed from a CUDA graph so the Python/ctypes call cost is not measured."""
    for i in range(8):
        fn(*sets[i % 8])
    torch.cuda.synchronize()
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(reps):
                fn(*sets[i % 8])
        run = g.replay
    else:
        def run():
            for i in range(reps):
                fn(*sets[i % 8])
    run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1000.0


for name, sets in (("uniform-random targets", rnd), ("smooth flow (grid + shift + noise)", smooth)):
    print(name)
    for label, fn in (
        ("rows kernel (row-major fp16 volume)", lambda a, b: ops.corr_lookup_pair(py_f, py_b, a, b, out[:Np], out[Np:])),
        ("tiled kernel, fp16x2 math", lambda a, b: ops.corr_lookup_tiled(tp_f, tp_b, a, b, out[:Np], out[Np:], math_f32=False)),
        ("tiled kernel, fp32 math", lambda a, b: ops.corr_lookup_tiled(tp_f, tp_b, a, b, out[:Np], out[Np:], math_f32=True)),
    ):
        us = timeit(fn, sets)
        print(f"  {label:40s} {us:8.1f} us   {alg / us / 1e3:8.1f} GB/s algorithmic")
us = timeit(lambda a, b: ops.corr_volume_pyramid_tiled(f[0], f[1]), rnd, 6, graph=False)
print(f"corr_volume_pyramid_tiled (3 pairs): {us:.1f} us")
us = timeit(lambda a, b: ops.corr_volume_pyramid(f[0], f[1], 4, torch.float16), rnd, 6, graph=False)
print(f"corr_volume_pyramid row-major fp16 (3 pairs): {us:.1f} us")
