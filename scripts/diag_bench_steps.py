"""Per-window device time of bench.py's FlowJob (C1) to see where the average comes from."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import bench
cfg = bench.CONFIGS["C1"]
job = bench.FlowJob(cfg, "cuda:0", "fp16", 32, 1, 0, 20)
wf, wn = job.clip(50000)
bench.timed_job(torch, None, job, 1, wf, wn, False)
fr, need = job.clip(0)
job.upload(fr, need)
torch.cuda.synchronize()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(21)]
p = job.proc
evs[0].record()
for i, f in enumerate(job.my_frames):
    job.out[job.slot[f]].copy_(p.compute_optical_flow_tiled(fr, f, return_device=True).permute(2, 0, 1))
    evs[i + 1].record()
torch.cuda.synchronize()
print("per-window ms:", [round(evs[i].elapsed_time(evs[i + 1]), 2) for i in range(20)])
print("stats", p.core.cache_stats)
ms, _ = bench.timed_job(torch, None, job, 1, *job.clip(7000), False)
print("timed_job again (new clip):", ms / 20, "ms/window")
s = bench.ClockSampler(0); s.start()
ms, _ = bench.timed_job(torch, None, job, 1, *job.clip(9000), False)
print("timed_job with the clock sampler running:", ms / 20, "ms/window", s.summary())
