"""MemFlow (MemFlowNet_skflow) online inference engine on libvfml kernels.

Mirrors InferenceCore.step (MemFlow/inference/inference_core_skflow.py:20-54) with the MemoryManager /
KeyValueMemoryStore state (memory_manager_skflow.py:38-124, kv_memory_store.py:32-91) kept in two device
buffers, and the network pieces of MemFlow.py:66-128,173-184 and sk2.py:86-109,182-215.

What is different from the reference and does not change results:
* q and k are constant within a frame, so the read-out affinity softmax(q.k^T * s) over the memory is built ONCE
  per frame ([P, T] matrix, T = (stored frames + 1) * P) instead of once per iteration
  (memory_manager_skflow.py:57-61 recomputes it 15 times); per iteration only `affinity @ values` runs, as a GEMM
  whose epilogue adds gamma * readout to the motion features (inference_core_skflow.py:43).
* The working memory is laid out [current frame | stored frames] so that the per-iteration value of the current
  frame lands at column 0; the softmax / weighted sum over memory elements does not depend on their order.
* The mask head runs on the last iteration only (its earlier outputs are discarded, :39-47).
* Activations live in pixel-major rows buffers; every torch.cat of the reference is a column range.

The update step reuses the MOF engine's PC-block driver, tiled correlation pyramid / lookup and CUDA-graph replay.
"""
import math
import os

import torch

from . import _lib
from . import ops as _ops
from .basic_encoder import BasicEncoder
from .mof import MOFEngine, _PCWeights, _pad8, _pad_act, _pad_k

ITERS = 15                                      # configs/sintel_memflownet.py:24 (val_decoder_depth)
TRAIN_AVG_LENGTH = (368 * 768 // 64) * 3 / 2    # configs/sintel_memflownet.py:28
DEFAULT_CONFIG = {"mem_every": 1, "max_mid_term_frames": 2, "min_mid_term_frames": 1,     # configs/sintel_memflownet.py:30-36
                  "train_avg_length": TRAIN_AVG_LENGTH, "val_decoder_depth": ITERS, "enable_long_term": False, "top_k": None}


class MemFlowEngine(MOFEngine):
    """One clip at a time: ``clear_memory()`` then ``step(images)`` per consecutive frame pair."""

    def __init__(self, state_dict, device, dtype=torch.float32, config=None, ops=None, encoder_dtype=None, use_graph=True,
                 tiled=None):
        cfg = dict(DEFAULT_CONFIG)
        cfg.update(config or {})
        if cfg.get("enable_long_term"):
            raise ValueError("enable_long_term is not supported (the reference asserts the same, memory_manager_skflow.py:30)")
        if cfg.get("top_k") is not None:
            raise NotImplementedError("top_k read-out is not built (configs ship top_k = None)")
        self.cfg = cfg
        self.arch = "memflow"
        self.ops = ops or _ops
        self.dev = torch.device(device)
        self.dt = dtype
        self.iters = int(cfg["val_decoder_depth"])
        self.mem_every = int(cfg["mem_every"])
        self.max_frames, self.min_frames = int(cfg["max_mid_term_frames"]), int(cfg["min_mid_term_frames"])
        self.max_stored = max(self.min_frames, self.max_frames - 1)       # frames alive after add_memory + compress
        self.tal = float(cfg["train_avg_length"])
        sd = {k[7:] if k.startswith("module.") else k: v for k, v in state_dict.items()}
        dev, dt = self.dev, dtype
        e = "update_block.encoder"
        self.tiled = (dtype == torch.float16) if tiled is None else bool(tiled)
        pack = self.ops.dwconv_pack_weights if (dtype == torch.float16 and hasattr(self.ops, "dwconv_pack_weights")) else None
        self.convc1 = _PCWeights(sd, e + ".convc1", 15, dev, dt, perm=self.ops.corr_lookup_perm() if self.tiled else None, pack=pack)
        self.convc2 = _PCWeights(sd, e + ".convc2", 15, dev, dt, pack=pack)
        self.convf2 = _PCWeights(sd, e + ".convf2", 15, dev, dt, pack=pack)
        self.conv = _PCWeights(sd, e + ".conv", 15, dev, dt, pack=pack)
        self.gru = _PCWeights(sd, "update_block.gru", 7, dev, dt, pack=pack)
        self.fhead = _PCWeights(sd, "update_block.flow_head", 15, dev, dt, pack=pack)
        # convf1 (2 -> 128, sk2.py:92) in the 4-column form of the shared glue kernel: its "forward" coordinates are a
        # constant pixel grid (flow 0), the real flow rides in the "backward" slot -> columns [0, 0, wx, wy]
        w2 = sd[e + ".convf1.weight"].to(dev).float().reshape(128, 2)
        self.convf1_w = torch.cat([torch.zeros_like(w2), w2], 1).contiguous()
        self.convf1_b = sd[e + ".convf1.bias"].to(dev).float().contiguous()
        self.w_v = _pad_k(sd["update_block.aggregator.to_v.weight"].to(dev).flatten(1).to(dt))
        self.gamma = float(sd["update_block.aggregator.gamma"].reshape(-1)[0])
        wqk = sd["att.to_qk.weight"].to(dev).flatten(1).to(dt)
        self.w_q, self.w_k = _pad_k(wqk[:128].contiguous()), _pad_k(wqk[128:].contiguous())
        m0 = sd["update_block.mask.0.weight"].to(dev)
        self.w_m0 = _pad_k(m0.permute(0, 2, 3, 1).reshape(256, 9 * 128).to(dt))
        self.b_m0 = sd["update_block.mask.0.bias"].to(dev).float().contiguous()
        self.w_m2 = _pad_k(sd["update_block.mask.2.weight"].to(dev).flatten(1).to(dt))
        self.mask_scale = 0.25                                            # sk2.py:214
        self.b_m2_x100 = (self.mask_scale * sd["update_block.mask.2.bias"].to(dev).float()).contiguous()
        enc_dt = encoder_dtype or torch.float32
        self.fnet = BasicEncoder(sd, "fnet", dev, "instance", enc_dt) if "fnet.conv1.weight" in sd else None
        self.cnet = BasicEncoder(sd, "cnet", dev, "batch", enc_dt) if "cnet.conv1.weight" in sd else None
        self._ws = {}
        self.use_graph = use_graph
        self.lookup_math_f32 = False
        self.graph_launches = 0
        self.use_side_stream = os.environ.get("VFML_NO_SIDE_STREAM") is None
        self.clear_memory()

    # ------------------------------------------------------------------ memory (MemoryManager + KeyValueMemoryStore)
    def clear_memory(self):
        """InferenceCore.clear_memory, inference_core_skflow.py:15-18"""
        self.curr_ti = -1
        self.last_mem_ti = -self.mem_every
        self.n_stored = 0            # frames in the working memory (rows P.. of ws['keys'], columns P.. of ws['vT'])

    def _add_memory(self, ws, P):
        """MemoryManager.add_memory + compress_features (memory_manager_skflow.py:99-124): append the frame's key and
        last value; at max_mid_term_frames keep the newest min_mid_term_frames.  Stored frames sit newest-first
        right behind the current-frame slot."""
        keys, vT = ws["keys"], ws["vT"]
        n = self.n_stored + 1
        if n >= self.max_frames:
            n = self.min_frames
        keep = n - 1                  # previously stored frames that survive (the newest ones = the first slots)
        if keep > 0:
            keys[2 * P:(2 + keep) * P] = keys[P:(1 + keep) * P].clone()
            vT[:, 2 * P:(2 + keep) * P] = vT[:, P:(1 + keep) * P].clone()
        keys[P:2 * P] = keys[0:P]
        vT[:, P:2 * P] = vT[:, 0:P]
        self.n_stored = n

    # ------------------------------------------------------------------ buffers
    def _workspace(self, Fc, h, w):
        key = (h, w)
        ws = self._ws.get(key)
        if ws is None:
            P = h * w
            Tmax = (1 + self.max_stored) * P
            ws = {
                "corr": self._rows_buf(P, 324),
                "cor_in": self._rows_buf(P, 256),
                "x256": self._rows_buf(P, 256),                # [cor 192 | flo 64]
                "x400": self._rows_buf(P, 400),                # scratch for the glue kernel's hidden-state slots (unused here)
                "f128": self._rows_buf(P, 128),
                "x512": self._rows_buf(P, 512),                # [net | inp | motion 126 + flow 2 | motion_global]
                "q": self._rows_buf(P, 128),
                "keys": self._rows_buf(Tmax, 128),             # [current | stored frames]
                "vT": torch.zeros((128, _pad8(Tmax)), dtype=self.dt, device=self.dev),
                "attn": torch.zeros((P, _pad_act(Tmax)), dtype=self.dt, device=self.dev),
                "logits": torch.empty((P, _pad_act(Tmax)), dtype=torch.float32, device=self.dev),
                "delta": self._rows_buf(P, 4, torch.float32),  # [delta x, delta y, 0, 0]
                "sa": self._rows_buf(P, 512),
                "sb": self._rows_buf(P, 512),
                "hid": self._rows_buf(P, 9 * 128),
                "hid2": self._rows_buf(P, 192), "sa2": self._rows_buf(P, 128), "sb2": self._rows_buf(P, 128),   # side stream (convf2)
                "mask": self._rows_buf(P, 576, torch.float32),
                "graphs": {},
            }
            # split-K partial sums of the read-out, sized for the worst memory length: the graphs captured per memory
            # length hold this buffer's address, so it is never replaced afterwards
            s_max = max(self._splits(P, k * P) for k in range(1, 2 + self.max_stored))
            ws["partial0"] = torch.empty(s_max * P * 128, dtype=torch.float32, device=self.dev)
            self._ws = {key: ws}
            self.n_stored = 0         # a new frame size invalidates the stored keys / values
        return ws

    # ------------------------------------------------------------------ pieces
    @torch.no_grad()
    def encode(self, images):
        """images [1, 2, 3, H, W] in [-1, 1] -> fmaps [2,256,h,w], net, inp [1,128,h,w]   (MemFlow.py:66-128)"""
        x = images.float()
        self._fork_side()                                   # the context encoder runs beside the feature encoder
        cnet = self.cnet(x[:, 0])
        self._back_to_main()
        fmaps = self.fnet(x[0])
        self._join_side()
        if cnet.is_cuda:
            cnet.record_stream(torch.cuda.current_stream(cnet.device))
        net, inp = torch.split(cnet, [128, 128], dim=1)
        return fmaps, torch.tanh(net).contiguous(), torch.relu(inp).contiguous()

    def _affinity(self, ws, P, T):
        """to_qk (MemFlow.py:110), then softmax over the T memory elements of q.k^T * 128^-1/2 * log_{train_avg_length} T
        (memory_manager_skflow.py:57-61, memory_util.py:57-61) -> ws['attn'][:, :T]"""
        o = self.ops
        inp = ws["x512"][:, 128:256]
        o.gemm(inp, self.w_q, ws["q"])
        o.gemm(inp, self.w_k, ws["keys"][0:P])
        scale = 128 ** -0.5 * math.log(T, self.tal)
        logits = self._view(ws["logits"], P, T)
        o.gemm(ws["q"], ws["keys"][0:T], logits, alpha=scale)
        o.softmax_rows(logits, ws["attn"][:, :T])

    def _update_step(self, ws, pyr, c1, grid, h, w, T):
        o = self.ops
        P = h * w
        x512, x256 = ws["x512"], ws["x256"]
        mf = x512[:, 256:384]
        full = slice(None)
        # flow branch (sk2.py:102-103) on a side stream beside the correlation branch (sk2.py:98-100)
        self._fork_side()
        o.motion_glue(grid, c1, ws["x400"], ws["f128"], mf, self.convf1_w, self.convf1_b, 1, h, w)   # convf1 + flow -> mf[:, 126:128]
        self._pcblock(ws["f128"], self.convf2, ws, 1, h, w, [(full, full, x256[:, 192:256], None)], scratch=("hid2", "sa2", "sb2"))
        self._back_to_main()
        if self.tiled:
            o.corr_lookup_tiled(pyr, None, c1, None, ws["corr"], None, math_f32=self.lookup_math_f32)
        else:
            o.corr_lookup(pyr, c1, 4, out=ws["corr"])
        self._pcblock(ws["corr"], self.convc1, ws, 1, h, w, [(full, full, ws["cor_in"], "gelu")])     # sk2.py:98
        self._pcblock(ws["cor_in"], self.convc2, ws, 1, h, w, [(full, full, x256[:, 0:192], None)])
        self._join_side()
        self._pcblock(x256, self.conv, ws, 1, h, w, [(full, full, x512[:, 256:382], None)])           # 126 + flow 2
        # value of the current frame (sk2.py:201-203), transposed straight into its memory slot; read-out + gamma
        vT = ws["vT"]
        o.gemm(self.w_v, mf, vT[:, 0:P])
        self._aggregate(ws, ws["attn"][:, :T], vT[:, :T], mf, x512[:, 384:512])
        self._pcblock(x512, self.gru, ws, 1, h, w, [(full, full, x512[:, 0:128], None)])
        self._pcblock(x512[:, 0:128], self.fhead, ws, 1, h, w, [(full, full, ws["delta"][:, 0:2], None)])
        o.coords_add(c1, grid, ws["delta"])

    def _body(self, ws, fmaps, net, inp, flow_init, h, w, n_stored):
        o = self.ops
        P = h * w
        T = (1 + n_stored) * P
        if self.tiled:
            pyr = o.corr_volume_pyramid_tiled(fmaps[0:1].contiguous(), fmaps[1:2].contiguous())
        else:
            pyr = o.corr_volume_pyramid(fmaps[0:1].contiguous(), fmaps[1:2].contiguous(), 4, self.dt)
        o.nchw_to_rows(net, ws["x512"][:, 0:128])
        o.nchw_to_rows(inp, ws["x512"][:, 128:256])
        self._affinity(ws, P, T)
        grid = o.coords_grid(1, h, w, self.dev)
        c1 = o.coords_grid(1, h, w, self.dev)
        c1 += flow_init                                   # MemFlow.py:95-96
        for _ in range(self.iters):
            self._update_step(ws, pyr, c1, grid, h, w, T)
        return c1, grid

    def _run(self, ws, fmaps, net, inp, flow_init, h, w):
        args = [t.float().contiguous() for t in (fmaps, net, inp, flow_init)]
        if not (self.use_graph and fmaps.is_cuda):
            return self._body(ws, *args, h, w, self.n_stored)
        key = (self.n_stored, self.iters)
        g = ws["graphs"].get(key)
        if g is None:
            static = [t.clone() for t in args]
            with _ops.gemm_autotuning():
                self._body(ws, *static, h, w, self.n_stored)            # eager warm-up; only the current-frame memory slot is written
            torch.cuda.synchronize(self.dev)
            graph = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with torch.cuda.graph(graph):
                coords = self._body(ws, *static, h, w, self.n_stored)
            g = ws["graphs"][key] = {"graph": graph, "static": static, "coords": coords, "launches": _lib.launch_count() - n0}
        for s_, t in zip(g["static"], args):
            s_.copy_(t)
        g["graph"].replay()
        self.graph_launches += g["launches"]
        return g["coords"]

    # ------------------------------------------------------------------ InferenceCore.step
    @torch.no_grad()
    def step(self, images, end=False, add_pe=False, flow_init=None, crop=None):
        """images [1, 2, 3, H, W] in [-1, 1], H and W multiples of 8 -> (flow_low [1,2,h,w], flow_up [1,2,H,W]).
        inference_core_skflow.py:20-54.  crop=(top, left, out_h, out_w) fuses InputPadder.unpad into the upsampling."""
        if images.dim() != 5 or images.shape[0] != 1 or images.shape[1] != 2 or images.shape[2] != 3:
            raise ValueError(f"step expects images [1, 2, 3, H, W], got {tuple(images.shape)}")
        if images.shape[-1] % 8 or images.shape[-2] % 8:
            raise ValueError("H and W must be multiples of 8 (pad with InputPadder first)")
        with self._guard():
            return self._step(images, end, flow_init, crop)

    def _step(self, images, end, flow_init, crop):
        o = self.ops
        self.curr_ti += 1
        is_mem_frame = (self.curr_ti - self.last_mem_ti >= self.mem_every) and not end
        fmaps, net, inp = self.encode(images)
        h, w = fmaps.shape[-2:]
        P = h * w
        ws = self._workspace(1, h, w)
        if flow_init is None:
            flow_init = torch.zeros((1, 2, h, w), dtype=torch.float32, device=self.dev)
        c1, grid = self._run(ws, fmaps, net, inp, flow_init.to(self.dev), h, w)
        flow_low = c1 - grid
        mk = self.mask_rows(ws, slice(0, 1), h, w, slice(0, 576))
        oh, ow = (crop[2], crop[3]) if crop else (8 * h, 8 * w)
        flow_up = torch.empty((1, 2, oh, ow), dtype=torch.float32, device=self.dev)
        o.convex_upsample8(c1.contiguous(), mk, out=flow_up, crop=crop, flow_is_coords=True)
        if is_mem_frame:
            self._add_memory(ws, P)
            self.last_mem_ti = self.curr_ti
        return flow_low, flow_up
