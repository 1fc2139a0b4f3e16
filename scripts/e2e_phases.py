"""Where the end-to-end call spends its time beyond the device step: phases of VideoFlowProcessor.compute_optical_flow."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
import numpy as np, torch
import cases
from oracle.mof import mof_synthetic_weights
from video_flow_ml_b200.processing import VideoFlowProcessor
p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp16")
p.core.load_state_dict(mof_synthetic_weights(0))
frames = cases.synthetic_clip(8, 436, 1024, seed=1234)
for i in range(3):
    p.compute_optical_flow(frames, i)
torch.cuda.synchronize()
acc = {}
def tick(name, t0):
    torch.cuda.synchronize(); t = time.perf_counter(); acc[name] = acc.get(name, 0.0) + (t - t0); return t
n = 8
for i in range(n):
    t = time.perf_counter()
    w = p.prepare_frame_sequence(frames, i % 8); t = tick("prepare (stack + H2D + convert)", t)
    f = p.core.compute_flow_from_tensor(w); t = tick("compute_flow_from_tensor", t)
    o = f.permute(1, 2, 0).cpu().numpy(); t = tick("permute + D2H + numpy", t)
t0 = time.perf_counter()
for i in range(n):
    p.compute_optical_flow(frames, i % 8)
torch.cuda.synchronize()
print("whole call: %.2f ms" % ((time.perf_counter() - t0) / n * 1e3))
for k, v in acc.items():
    print(f"  {k:40s} {v / n * 1e3:7.2f} ms (synchronised after each phase)")
