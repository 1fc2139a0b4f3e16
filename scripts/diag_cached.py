"""Where does the cached sliding-window path spend its time?  (GPU box diagnostic)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch, cases
from oracle.mof import mof_synthetic_weights
from video_flow_ml_b200.processing import VideoFlowProcessor
p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp16")
p.core.load_state_dict(mof_synthetic_weights(0))
p.device_frame_capacity = 64
def clip(o): return cases.sliding_clip(o, o + 24, 436, 1024)
fr = clip(0)
for i in range(2, 8): p.compute_optical_flow(fr, i, return_device=True)
torch.cuda.synchronize()
for mode in ("nosync", "sync", "nosync"):
    fr = clip(1000 * (1 + hash(mode) % 7) + int(time.time()) % 100)
    for f in fr: p._device_frame(f, p._frame_key(f))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(); cpu = 0.0
    for i in range(2, 22):
        c0 = time.perf_counter()
        out = p.compute_optical_flow(fr, i, return_device=True)
        cpu += time.perf_counter() - c0
        if mode == "sync": torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    print(f"{mode}: {e0.elapsed_time(e1) / 20:.2f} ms/window on the device, CPU enqueue {1000 * cpu / 20:.2f} ms/window, wall {(time.perf_counter() - t0) * 50:.2f} ms/window", flush=True)
# phases of one steady-state window (synchronised between phases)
import numpy as np
fr = clip(777)
for i in range(2, 6): p.compute_optical_flow(fr, i, return_device=True)
torch.cuda.synchronize()
core = p.core
def timed(fn):
    torch.cuda.synchronize(); t = time.perf_counter(); r = fn(); torch.cuda.synchronize(); return r, (time.perf_counter() - t) * 1e3
_, t_key = timed(lambda: [p._frame_key(f) for f in fr[4:9]])
dev = p._device_frame(fr[9], p._frame_key(fr[9]))
x = p._chw(dev)[None]
x = torch.nn.functional.pad(x, [0, 0, 2, 2], mode="replicate")
_, t_f = timed(lambda: core.model.encode_frames(x))
_, t_c = timed(lambda: core.model.encode_frames(x, context=True))
x5 = torch.cat([x] * 5)
_, t_f5 = timed(lambda: core.model.encode_frames(x5))
print(f"keys {t_key:.2f} ms; fnet(1 frame) {t_f:.2f} ms; cnet(1 frame) {t_c:.2f} ms; fnet(5 frames) {t_f5:.2f} ms")
