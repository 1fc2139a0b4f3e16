#!/usr/bin/env python
"""Benchmark of the MOF optical-flow hot path (BASELINE.json: MOF seq-5 436x1024 flow frames/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--precision fp32|fp16] [--impl reference]

A "step" is one window (= one output flow frame = one call of compute_optical_flow): encoders on the
window's frames, two correlation pyramids, GMA attention, 32 update iterations, mask head + convex
upsampling of the consumed flow.  Synthetic clip (seed 1234) and seeded synthetic weights: there are
no datasets or checkpoints offline.

value : windows/s with the window tensor already resident in HBM (device timed, CUDA events).
e2e   : windows/s through VideoFlowProcessor.compute_optical_flow with HOST numpy frames
        (u8 upload of the window and fp32 download of the flow inside the timed region).
roofline : the correlation lookup kernel (HBM bound), algorithmic bytes / CUDA-event time.
cpu_baseline : the oracle (CPU port of the reference) on this box's host cores, bounded sample.

N > 1 (torchrun): windows are sharded over ranks, no data-path collective; the finished flows are
gathered to rank 0 with NCCL inside the timed region; value = all ranks' windows / max-rank time.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

H, W, T, ITERS, CLIP = 436, 1024, 5, 32, 8
WORKLOAD = "MOF_sintel T=5, 8 synthetic 436x1024 frames, 32 iterations (BASELINE.json configs[0]/[1] shape C1)"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_evt = index, [], threading.Event()

    def run(self):
        while not self.stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop_evt.wait(0.2)

    def summary(self):
        self.stop_evt.set()
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        reasons = []
        for i, name in enumerate(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), start=2):
            if any(len(r) > i and r[i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        mx = max((int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


def cpu_port_times(n_iter=2):
    """Oracle (CPU restatement of the reference) on the C1 window: returns (t_encode, t_setup, t_iter, cores)."""
    import torch
    import cases
    from oracle import mof as omof
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = omof.mof_synthetic_weights(0)
    frames = cases.synthetic_clip(T, H, W, seed=1234)
    win = omof.pad_frames(omof.frames_to_window(frames, T // 2, T))
    with torch.no_grad():
        t0 = time.perf_counter()
        fmaps, net, inp = omof.mof_encode(win, sd)
        t_enc = time.perf_counter() - t0
        t0 = time.perf_counter()
        omof.mof_refine(fmaps, net, inp, sd, iters=1)
        t1 = time.perf_counter() - t0
        t0 = time.perf_counter()
        omof.mof_refine(fmaps, net, inp, sd, iters=1 + n_iter)
        t2 = time.perf_counter() - t0
    t_iter = max((t2 - t1) / n_iter, 1e-9)
    return t_enc, max(t1 - t_iter, 0.0), t_iter, cores


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm (oracle port) on the host cores, all threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for _ in range(min(max(args.steps, 1), 3)):      # bounded: at most 3 full-size samples (~20 s each)
        t_enc, t_setup, t_iter, cores = cpu_port_times(1)
        vals.append(t_enc + t_setup + ITERS * t_iter)
    t = sum(vals) / len(vals)
    sample = ("per step: oracle MOF window 436x1024 T=5 at full size: encoders + 2 corr pyramids + attention + 2 measured "
              "update iterations, extrapolated linearly to 32 iterations")
    v = 1.0 / t
    print(json.dumps({"impl": "reference", "metric": "MOF seq5 436x1024 flow frames/s", "value": v, "unit": "frames/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000 * t,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": WORKLOAD},
                      "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--precision", default=os.environ.get("VFML_PRECISION", "fp16"))
    ap.add_argument("--iters", type=int, default=ITERS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import cases
    from video_flow_ml_b200 import _lib, ops, sharding
    from video_flow_ml_b200.processing import VideoFlowProcessor
    from oracle.mof import mof_synthetic_weights      # seeded weight generator only (a deterministic RNG)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    warm = max(args.warmup, 3)

    proc = VideoFlowProcessor(dev, sequence_length=T, precision=args.precision)
    proc.core.load_state_dict(mof_synthetic_weights(0), iters=args.iters)
    # every rank owns its own shard of windows: a different clip per rank (weak scaling)
    frames = cases.synthetic_clip(CLIP, H, W, seed=1234 + rank)
    n_steps = args.steps
    windows = [proc.prepare_frame_sequence(frames, i % CLIP) for i in range(CLIP)]
    out = torch.empty((n_steps, 2, H, W), dtype=torch.float32, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident arm ("value")
    for i in range(warm):
        proc.core.compute_flow_from_tensor(windows[i % CLIP], out=out[i % n_steps])
    if world > 1:      # untimed: NCCL sets up its point-to-point channels lazily, on the first gather
        sharding.gather_flows(out, world * n_steps, world, rank, dst=0)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = _lib.launch_count() + proc.core.model.graph_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(n_steps):
        proc.core.compute_flow_from_tensor(windows[i % CLIP], out=out[i])
    if world > 1:      # the only communication of the job: finished flows to rank 0 (NCCL over NVLink)
        sharding.gather_flows(out, world * n_steps, world, rank, dst=0)
    e1.record()
    barrier()
    launches = _lib.launch_count() + proc.core.model.graph_launches - n0     # direct launches + kernels replayed from the CUDA graph
    ms = e0.elapsed_time(e1)
    clocks = sampler.summary()

    # ---- end-to-end arm through the public API with host frames
    for i in range(2):
        proc.compute_optical_flow(frames, i)
    barrier()
    t0 = time.perf_counter()
    for i in range(n_steps):
        flow_np = proc.compute_optical_flow(frames, i % CLIP)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    h2d = T * H * W * 3                  # u8 window upload
    d2h = int(flow_np.nbytes)

    # ---- roofline of the lookup kernel, timed live with CUDA events on the launching stream
    eng = proc.core.model
    hp, wp = (H + 7) // 8 * 8 // 8, W // 8
    Fc, P = T - 2, hp * wp
    fmaps, net, inp = eng.encode(torch.nn.functional.pad(windows[0], [0, 0, 2, 2, 0, 0], mode="replicate"))
    fpyr, bpyr = eng.build_pyramids(fmaps)
    g = torch.Generator(device="cpu").manual_seed(7)
    # 8 coordinate sets with uniformly random in-range targets: successive launches touch different texels of
    # the 1.58 GB of volumes (8 x ~100 MB of sectors >> 126 MB L2), so every launch streams from HBM
    n_sets = 8
    lo = torch.tensor([4.0, 4.0]).view(1, 2, 1, 1)
    span = torch.tensor([wp - 9.0, hp - 9.0]).view(1, 2, 1, 1)
    csets = [((lo + span * torch.rand((Fc, 2, hp, wp), generator=g)).to(dev).contiguous(),
              (lo + span * torch.rand((Fc, 2, hp, wp), generator=g)).to(dev).contiguous()) for _ in range(n_sets)]
    lk_out = torch.zeros((2 * Fc * P, 328), dtype=eng.dt, device=dev)[:, :324]
    reps = 48
    if eng.tiled:
        lk_kernel = "corr_lookup_tiled_kernel"
        lookup = lambda cf, cb: ops.corr_lookup_tiled(fpyr, bpyr, cf, cb, lk_out[:Fc * P], lk_out[Fc * P:], math_f32=eng.lookup_math_f32)
    else:
        lk_kernel = "corr_lookup_rows_kernel"
        lookup = lambda cf, cb: ops.corr_lookup_pair(fpyr, bpyr, cf, cb, lk_out[:Fc * P], lk_out[Fc * P:])
    for i in range(n_sets):
        lookup(*csets[i])
    torch.cuda.synchronize()
    # the launches are replayed from a CUDA graph (as in the engine) so that the ~20 us ctypes call is not what is timed
    lk_graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(lk_graph):
        for i in range(reps):
            lookup(*csets[i % n_sets])
    lk_graph.replay()
    l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    l0.record()
    lk_graph.replay()
    l1.record()
    torch.cuda.synchronize()
    lk_ms = l0.elapsed_time(l1) / reps          # one launch = forward + backward lookup of one iteration
    es = 4 if eng.dt == torch.float32 else 2
    # DRAM bytes per launch of the roofline kernel on this workload, from the committed ncu --set full capture
    NCU_TRAFFIC = {"corr_lookup_tiled_kernel": 74438912 + 3937792}
    alg_bytes = 2 * Fc * P * (4 * 100 * es + 8 + 324 * es)   # per pixel-direction 100 texels/level read + coords + 324 written
    peak, peak_kind = peaks()
    achieved = alg_bytes / (lk_ms * 1e-3) / 1e9
    del fpyr, bpyr

    # ---- second roofline, for the kernel that dominates the step by time (gemm_tc_kernel, ~65 %): the 30 pointwise-
    # convolution GEMMs of one update iteration on synthetic activations of the real shapes, replayed from a CUDA graph
    gemm_roof = None
    if eng.dt == torch.float16:
        Np = Fc * P
        rb = lambda n, c: torch.randn((n, (c + 7) // 8 * 8), device=dev).to(eng.dt)[:, :c]
        calls, flops = [], 0.0
        for M, C, Co in ((2 * Np, 324, 128), (Np, 256, 192), (Np, 128, 64), (Np, 400, 172), (Np, 512, 128), (Np, 128, 4)):
            hid = int(1.5 * C)
            x, t1, x2, x3, o_ = rb(M, C), rb(M, hid), rb(M, C), rb(M, C), rb(M, Co)
            w1a, w1b, wpw, w2a, w2b = rb(hid, C), rb(C, hid), rb(C, C), rb(hid, C), rb(Co, hid)
            bh, bc, bo = (torch.randn(n, device=dev) for n in (hid, C, Co))
            calls += [lambda x=x, w=w1a, o=t1, b=bh: ops.gemm(x, w, o, bias=b, act="gelu", weights=True),
                      lambda x=t1, w=w1b, o=x2, b=bc, r=x: ops.gemm(x, w, o, bias=b, res=r, act="gelu", dw=(b, b), weights=True),
                      lambda x=x3, w=wpw, o=x2, b=bc: ops.gemm(x, w, o, bias=b, res=x, act="gelu", weights=True),
                      lambda x=x2, w=w2a, o=t1, b=bh: ops.gemm(x, w, o, bias=b, act="gelu", weights=True),
                      lambda x=t1, w=w2b, o=o_, b=bo: ops.gemm(x, w, o, bias=b, weights=True)]
            flops += 2.0 * M * (3 * hid * C + C * C + hid * Co)      # ffn1.0, ffn1.2, ffn2.0 (hid x C each), pw (C x C), ffn2.2 (hid x Co)
        for c_ in calls:
            c_()
        torch.cuda.synchronize()
        gg = torch.cuda.CUDAGraph()
        greps = 4
        with torch.cuda.graph(gg):
            for _ in range(greps):
                for c_ in calls:
                    c_()
        gg.replay()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        g0.record()
        gg.replay()
        g1.record()
        torch.cuda.synchronize()
        g_ms = g0.elapsed_time(g1) / greps
        tpeak = None
        try:
            tpeak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        except Exception:
            pass
        ach = flops / (g_ms * 1e-3) / 1e12
        gemm_roof = {"kernel": "gemm_tc_kernel", "bound": "tensor", "achieved": ach, "peak": tpeak, "unit": "TFLOP/s",
                     "frac": ach / tpeak if tpeak else None, "traffic": None, "launches": len(calls),
                     "flops_per_iteration": flops, "us_per_iteration": g_ms * 1000.0,
                     "note": "the 30 pointwise-convolution GEMMs of one update iteration (SURVEY 8d: 114.4 GMAC minus the "
                             "depthwise and mask-head terms), fused epilogues included, timed alone by graph replay; peak = "
                             "measured dense bf16 burst figure"}
        del calls

    tmax = torch.tensor([ms, e2e_s * 1000.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(tmax[0]), float(tmax[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    total = world * n_steps
    line = {
        "metric": "MOF seq5 436x1024 flow frames/s", "value": total / (ms_max / 1000.0), "unit": "frames/s",
        "n_gpus": world, "steps": n_steps, "warmup": warm, "ms_per_step": ms_max / n_steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.precision == "fp32" else "f16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "iterations": args.iters, "precision": args.precision,
                   "l2": "inputs larger than L2 (%s of correlation volumes per window)" % ("1.58 GB" if args.precision == "fp32" else "811 MB"),
                   "sharding": f"{world} rank(s), windows sharded, flows gathered to rank 0 (NCCL) inside the timed region"},
        "clocks": clocks,
        "e2e": {"value": total / (e2e_ms_max / 1000.0), "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "roofline": {"kernel": lk_kernel, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": NCU_TRAFFIC.get(lk_kernel), "peak_source": peak_kind,
                     "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full capture of this command "
                                       "(profiles/r1_lookup_tiled_ncu_summary.txt; cold L2)" if lk_kernel in NCU_TRAFFIC else None,
                     "algorithmic_bytes_per_launch": alg_bytes, "us_per_launch": lk_ms * 1000.0},
    }
    if gemm_roof is not None:
        line["roofline_gemm"] = gemm_roof
    if not args.no_cpu_baseline and world == 1:
        t_enc, t_setup, t_iter, cores = cpu_port_times(2)
        tw = t_enc + t_setup + ITERS * t_iter
        line["cpu_baseline"] = {"value": 1.0 / tw, "unit": "frames/s", "cores": cores, "kind": "port",
                                "sample": "oracle MOF window 436x1024 T=5 at full size: encoders + corr pyramids + attention + 2 "
                                          "measured update iterations, extrapolated linearly to 32"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
