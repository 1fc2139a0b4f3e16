"""What the host-buffer call adds to the device-resident step (C1, fp16 tier): per-call wall time of compute_optical_flow with
numpy out / with return_device=True + synchronize, and the device->host tail done three ways."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch
import cases
from oracle.mof import mof_synthetic_weights
from video_flow_ml_b200.processing import VideoFlowProcessor

p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp16")
p.core.load_state_dict(mof_synthetic_weights(0))
frames = cases.sliding_clip(0, 60, 436, 1024, seed=7)
for i in range(12):
    p.compute_optical_flow(frames, i)
torch.cuda.synchronize()

def wall(fn, idx):
    ts = []
    for i in idx:
        t0 = time.perf_counter(); fn(i); ts.append((time.perf_counter() - t0) * 1e3)
    return np.median(ts), ts

m, ts = wall(lambda i: p.compute_optical_flow(frames, i), range(12, 28))
print(f"numpy out                       : median {m:.2f} ms per call   {[round(x, 2) for x in ts]}")
def dev(i):
    p.compute_optical_flow(frames, i, return_device=True); torch.cuda.synchronize()
m, ts = wall(dev, range(28, 44))
print(f"return_device=True + synchronize: median {m:.2f} ms per call   {[round(x, 2) for x in ts]}")
def dev_nosync(i):
    p.compute_optical_flow(frames, i, return_device=True)
t0 = time.perf_counter()
for i in range(44, 60):
    dev_nosync(i)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"return_device=True, no sync     : host {(t1 - t0) / 16 * 1e3:.2f} ms per call, device-bound {(t2 - t0) / 16 * 1e3:.2f} ms per call")

flow = torch.randn(2, 436, 1024, device="cuda")
pin = torch.empty((436, 1024, 2), dtype=torch.float32, pin_memory=True)
def a(): return flow.permute(1, 2, 0).cpu().numpy()
def b():
    pin.copy_(flow.permute(1, 2, 0), non_blocking=True); torch.cuda.current_stream().synchronize(); return pin.numpy().copy()
def c():
    t = flow.permute(1, 2, 0).contiguous(); pin.copy_(t, non_blocking=True); torch.cuda.current_stream().synchronize(); return pin.numpy().copy()
for name, fn in (("permute.cpu().numpy()", a), ("pinned copy_ of the permuted view + np copy", b), ("contiguous + pinned copy_ + np copy", c)):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20): fn()
    print(f"D2H tail, {name:45s}: {(time.perf_counter() - t0) / 20 * 1e3:.3f} ms")
f8 = frames[0]
t0 = time.perf_counter()
for _ in range(50): VideoFlowProcessor._frame_key(f8)
print(f"_frame_key: {(time.perf_counter() - t0) / 50 * 1e3:.3f} ms per frame")
