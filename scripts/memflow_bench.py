"""MemFlow online throughput on one GPU at the C1 frame size (436 x 1024, 15 iterations, one stored memory frame):
frames/s device-resident (frame pairs already in HBM) and end to end through MemFlowProcessor (host uint8 frames in,
numpy flow out).  Secondary measurement for DESIGN.md; bench.py stays on BASELINE.json's MOF metric."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from oracle.memflow import memflow_synthetic_weights      # seeded weight generator only
from video_flow_ml_b200 import _lib
from video_flow_ml_b200.processing import InferenceCore, MemFlowProcessor

precision = sys.argv[1] if len(sys.argv) > 1 else "fp16"
H, W, n_frames = 436, 1024, 12
rng = np.random.default_rng(1234)
base = rng.integers(0, 256, (H + 64, W + 64, 3), dtype=np.uint8)
frames = [np.ascontiguousarray(base[8 + 2 * i:8 + 2 * i + H, 8 + 3 * i:8 + 3 * i + W]) for i in range(n_frames)]
p = MemFlowProcessor("cuda", precision=precision)
p.core_engine.load_state_dict(memflow_synthetic_weights(0))
eng = p.core_engine.model

x = torch.stack([torch.from_numpy(f).permute(2, 0, 1).float() for f in frames])[None].cuda()
x = torch.nn.functional.pad(2 * (x / 255.0) - 1.0, [0, 0, 2, 2, 0, 0], mode="replicate")    # 436 -> 440
ic = InferenceCore(eng)
for t in range(3):
    ic.step(x[:, t:t + 2])
torch.cuda.synchronize()
n0 = _lib.launch_count() + eng.graph_launches
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ic.clear_memory()
ic.step(x[:, 0:2])
e0.record()
steps = n_frames - 2
for t in range(1, 1 + steps):
    ic.step(x[:, t:t + 2])
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
launches = (_lib.launch_count() + eng.graph_launches - n0) / (steps + 1)
print(f"memflow {precision} online (memory engaged), device-resident: {ms:.2f} ms / frame = {1000 / ms:.1f} frames/s, "
      f"{launches:.0f} vfml launches / frame")

p.compute_optical_flow(frames, 3)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(2, 2 + 8):
    p.compute_optical_flow(frames, i)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 8 * 1e3
print(f"memflow {precision} MemFlowProcessor.compute_optical_flow (host frames in, numpy out, fresh memory per call): "
      f"{ms:.2f} ms / frame = {1000 / ms:.1f} frames/s")
