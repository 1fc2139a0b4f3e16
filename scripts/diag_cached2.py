import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch, cases
from oracle.mof import mof_synthetic_weights
from video_flow_ml_b200.processing import VideoFlowProcessor
p = VideoFlowProcessor("cuda", sequence_length=5, precision="fp16")
p.core.load_state_dict(mof_synthetic_weights(0))
p.device_frame_capacity = 64
eng = p.core.model
fr = cases.sliding_clip(0, 30, 436, 1024)
for i in range(2, 8): p.compute_optical_flow(fr, i, return_device=True)
torch.cuda.synchronize()
def ev(fn, n=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
x = torch.nn.functional.pad(p._chw(p._device_frame(fr[9], p._frame_key(fr[9])))[None], [0, 0, 2, 2], mode="replicate")
print("encode_new_frames graph (1 fnet + 1 cnet):", ev(lambda: eng.encode_new_frames(x, x)), "ms")
eng.use_graph = False
print("encode_new_frames eager two streams:", ev(lambda: eng.encode_new_frames(x, x)), "ms")
print("fnet only eager:", ev(lambda: eng.encode_new_frames(x, None)), "ms;  cnet only eager:", ev(lambda: eng.encode_new_frames(None, x)), "ms")
eng.use_graph = True
fm, ctx = eng.encode_new_frames(x, x)
fmaps = torch.cat([fm] * 5); net = torch.cat([ctx[0]] * 3); inp = torch.cat([ctx[1]] * 3)
out = torch.empty((2, 436, 1024), device="cuda")
print("refine (graph replay + mask + upsample):", ev(lambda: eng.refine(fmaps, net, inp, select=3, crop=(2, 0, 436, 1024), out=out)), "ms")
ws = eng._workspace(3, 55, 128)
g = ws["graphs"][(False, 0)]["graph"]
print("iteration graph replay alone:", ev(lambda: g.replay()), "ms")
win = p.prepare_frame_sequence(fr, 12)
print("uncached compute_flow_from_tensor:", ev(lambda: p.core.compute_flow_from_tensor(win, out=out)), "ms")
xw = torch.nn.functional.pad(win, [0, 0, 2, 2, 0, 0], mode="replicate")
print("uncached encode (5 + 3, two streams, eager):", ev(lambda: eng.encode(xw)), "ms")
i = [12]
def step():
    i[0] += 1
    p.compute_optical_flow(fr, i[0], return_device=True)
print("cached sliding step:", ev(step, n=12), "ms")
