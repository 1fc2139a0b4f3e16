import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch, cases
from oracle import mof as omof
from video_flow_ml_b200.processing import VideoFlowProcessor
g = np.load(os.path.join(ROOT, "tests/golden/mof_tiled.npz"))
sd = omof.mof_synthetic_weights(0, gain=1.0)
frames = [f.astype(np.float32) for f in cases.synthetic_clip(5, 288, 448, seed=5)]
def epe(a, b): return float(np.sqrt(((a - b) ** 2).sum(-1)).max())
for prec in ("fp16", "fp32"):
    for graph in (True, False):
        p = VideoFlowProcessor("cuda", sequence_length=5, precision=prec)
        p.core.load_state_dict(sd, iters=4)
        p.core.model.use_graph = graph
        orig = p.calculate_tile_grid
        p.calculate_tile_grid = lambda w, h, tile_size=160: orig(w, h, tile_size)
        p.set_tile_mode(True)
        for cache in (True, False):
            p.set_feature_cache(cache)
            for order in ((2, 0), (0, 2)):
                p.core.clear_feature_cache()
                for idx in order:
                    full = p.compute_optical_flow_tiled(list(frames), idx)
                    ref = g["tiled_f32_2"] if idx == 2 else None
                    errs = []
                    for ti in orig(448, 288, 160)[4]:
                        x, y, w, h = ti["x"], ti["y"], ti["width"], ti["height"]
                        if idx == 2:
                            errs.append(epe(full[y:y + h, x:x + w], ref[y:y + h, x:x + w]))
                    if idx == 0:
                        sub = full[::3, ::3]; d = np.sqrt(((sub - g["tiled_f32_0"]) ** 2).sum(-1))
                        yy, xx = np.unravel_index(d.argmax(), d.shape)
                        errs = [float(d.max()), f"at y={3 * yy} x={3 * xx}"]
                    print(prec, "graph" if graph else "eager", "cache" if cache else "nocache", "order", order, "idx", idx, errs, flush=True)
        del p
