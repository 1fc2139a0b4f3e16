"""Volume + pyramid build: fused epilogue (corr_fused.cu) against GEMM -> row-major level 0 -> corr_tile_pyramid_kernel.
Same process, interleaved, CUDA events on the launching stream; buffers larger than L2 are rewritten between calls by the build itself
(a 2-pair pyramid is 270 MB).   python scripts/micro/corr_fused_ab.py > profiles/r2_corr_fused_ab.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from video_flow_ml_b200 import ops

torch.manual_seed(0)
dev = torch.device("cuda:0")
for name, n, C, h, w in (("C1 steady state (2 x 1 new pair)", 1, 256, 55, 128), ("C1 cold window (2 x 3 pairs)", 3, 256, 55, 128),
                         ("C3 (T=9: 2 x 1 new pair)", 1, 256, 47, 156), ("4K tile 1280x1280 (2 x 1 pair)", 1, 256, 160, 160),
                         ("C = 128, 2 pairs at C1 size", 2, 128, 55, 128)):
    f1 = torch.randn(n, C, h, w, device=dev)
    f2 = torch.randn(n, C, h, w, device=dev)
    out = ops.TiledPyramid(ops._tiled_levels(n * h * w, h, w, dev), n, h, w)
    ws = torch.empty(ops._lib.load().vfml_corr_volume_tiled_workspace(n, C, h, w), dtype=torch.uint8, device=dev)
    res = {}
    for rep in range(3):
        for mode in ("1", "0"):
            os.environ["VFML_CORR_FUSED"] = mode
            for _ in range(2):
                ops.corr_volume_pyramid_tiled(f1, f2, workspace=ws, out=out)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                ops.corr_volume_pyramid_tiled(f1, f2, workspace=ws, out=out)
            e1.record()
            torch.cuda.synchronize()
            res.setdefault(mode, []).append(e0.elapsed_time(e1) / 5 * 1e3)
    P = h * w
    pyr_bytes = sum(l.numel() for l in out.levels)
    fu, un = min(res["1"]), min(res["0"])
    print(f"{name}: P = {P}, pyramid {pyr_bytes / 1e6:.0f} MB | fused {fu:.0f} us ({pyr_bytes / fu / 1e6:.2f} TB/s of pyramid bytes) | "
          f"unfused {un:.0f} us | x{un / fu:.2f}   (all runs: fused {[round(x) for x in res['1']]}, unfused {[round(x) for x in res['0']]})", flush=True)
