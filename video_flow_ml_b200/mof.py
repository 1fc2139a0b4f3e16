"""VideoFlow MOF (multi-frame) inference engine on libvfml kernels.

Mirrors MOFNet.forward (VideoFlow/core/Networks/MOFNetStack/network.py:138-215) for eval:
encoders -> two all-pairs correlation pyramids -> GMA attention -> 32 x {2 lookups, motion
encoder, aggregate, PC-block "GRU", flow head} -> mask head + convex 8x upsampling.

Differences from the reference that do not change results: the mask head and the
upsampling run on the last iteration only, and ``select`` lets the caller upsample only the
flow it consumes.  All activations of the update step live in pixel-major "rows" buffers
(NHWC) so that every 1x1 convolution is a plain K-contiguous GEMM and every ``torch.cat`` of
the reference is just a column range of a wider buffer.

Weights use the upstream checkpoint key names (``update_block.encoder.convc1.ffn1.0.weight``,
``att.to_qk.weight`` ...).  ``dtype`` selects the tier: float32 (exact fp32 FFMA kernels,
<= 1e-2 px vs the fp32 reference) or float16 (tensor-core tier; the reference itself runs
this part under fp16 autocast, network.py:149-192).
"""
import os

import torch

from . import _lib
from . import ops as _ops
from .twins import TwinsStage01

ITERS = 32          # VideoFlow/configs/multiframes_sintel_submission.py:36 (decoder_depth)
CORR_LEVELS = 4
CORR_RADIUS = 4     # forced, network.py:23


def _pad8(c):
    return (c + 7) // 8 * 8


_SMS = {}


def _sm_count():
    """SM count of the current CUDA device (148 on B200; also the answer without a device, for the CPU data-flow tests)."""
    if not torch.cuda.is_available():
        return 148
    d = torch.cuda.current_device()
    if d not in _SMS:
        _SMS[d] = torch.cuda.get_device_properties(d).multi_processor_count
    return _SMS[d]


def _pad_act(c):
    """Row pitch of activation buffers: a multiple of 16 elements, so that every row of an fp16 buffer starts on a 32-byte
    sector and the GEMM epilogue's 16-column pieces are single 256-bit stores (gemm_tc.cu)."""
    return (c + 15) // 16 * 16


def _pad_k(w):
    """[N, K] weight as a rows view whose pitch is a multiple of 8 elements (16-byte rows: TMA requirement)."""
    n, k = w.shape
    buf = torch.zeros((n, _pad8(k)), dtype=w.dtype, device=w.device)
    buf[:, :k] = w
    return buf[:, :k]


class _PCWeights:
    """PCBlock4_Deep_nopool_res parameters (stack.py:8-32) prepacked for the kernels."""

    def __init__(self, sd, p, k, dev, dt, perm=None, pack=None):
        """perm (optional, LongTensor [Cin]): the block's input arrives with channel `new` holding the
        reference's channel perm[new]; every op of the block up to ffn2 is channel-wise or a 1x1
        convolution, so the permutation is folded into the weights once and costs nothing at run time."""
        g = lambda n: sd[f"{p}.{n}"].to(dev)
        ident = lambda t: t
        pin = ident if perm is None else (lambda t: t.index_select(t.dim() - 1, perm.to(t.device)))   # input-channel axis (last)
        pout = ident if perm is None else (lambda t: t.index_select(0, perm.to(t.device)))            # output-channel axis (first)
        mat = lambda n, f=ident: _pad_k(f(g(n + ".weight").flatten(1)).to(dt))
        vec = lambda n, f=ident: f(g(n + ".bias").float()).contiguous()
        self.k = k
        self.w1a, self.b1a = mat("ffn1.0", pin), vec("ffn1.0")
        self.w1b, self.b1b = mat("ffn1.2", pout), vec("ffn1.2", pout)
        self.dw1_w = pout(g("conv_list.0.weight").float().flatten()).contiguous()      # k = 1 depthwise
        self.dw1_b = vec("conv_list.0", pout)
        C = self.dw1_w.numel()
        self.dwk_w = pout(g("conv_list.1.weight").float().reshape(C, k * k)).t().contiguous()   # tap-major [k*k, C]
        self.dwk_b = vec("conv_list.1", pout)
        self.wpw, self.bpw = mat("pw", lambda t: pin(pout(t))), vec("pw", pout)
        self.w2a, self.b2a = mat("ffn2.0", pin), vec("ffn2.0")
        self.w2b, self.b2b = mat("ffn2.2"), vec("ffn2.2")
        self.cin, self.hid, self.cout = C, self.w1a.shape[0], self.w2b.shape[0]
        # tensor-core depthwise (fp16 tier): banded-Toeplitz fragment table, packed once
        self.dwk_packed = pack(self.dwk_w, k) if pack is not None else None


class MOFEngine:
    """Stateless-per-window MOF refinement; buffers are cached per (F, h, w)."""

    def __init__(self, state_dict, device, dtype=torch.float32, iters=ITERS, ops=None, encoder_dtype=None, use_graph=True,
                 tiled=None, arch="mof"):
        """arch='mof': MOFNetStack (multi-frame, temporal hidden state).  arch='bof': BOFNet (3 frames, no hidden
        state, conv block 256 -> 124, mask scale 0.25, mask channels ordered [2->1 | 2->3]),
        BOFNet/network.py:82-145 + BOFNet/sk2.py:33-92."""
        if arch not in ("mof", "bof"):
            raise ValueError(f"arch must be 'mof' or 'bof', got {arch}")
        self.arch = arch
        self.ops = ops or _ops
        self.dev = torch.device(device)
        self.dt = dtype
        self.iters = iters
        sd = {k[7:] if k.startswith("module.") else k: v for k, v in state_dict.items()}
        dev, dt = self.dev, dtype
        e = "update_block.encoder"
        # fp16 tier: the correlation features come from the tiled lookup in its permuted channel order
        self.tiled = (dtype == torch.float16) if tiled is None else bool(tiled)
        pack = self.ops.dwconv_pack_weights if (dtype == torch.float16 and hasattr(self.ops, "dwconv_pack_weights")) else None
        self.convc1 = _PCWeights(sd, e + ".convc1", 15, dev, dt, perm=self.ops.corr_lookup_perm() if self.tiled else None, pack=pack)
        self.convc2 = _PCWeights(sd, e + ".convc2", 15, dev, dt, pack=pack)
        self.convf2 = _PCWeights(sd, e + ".convf2", 15, dev, dt, pack=pack)
        self.conv = _PCWeights(sd, e + ".conv", 15, dev, dt, pack=pack)
        self.gru = _PCWeights(sd, "update_block.gru", 7, dev, dt, pack=pack)
        self.fhead = _PCWeights(sd, "update_block.flow_head", 15, dev, dt, pack=pack)
        self.convf1_w = sd[e + ".convf1_.weight"].to(dev).float().reshape(128, 4).contiguous()
        self.convf1_b = sd[e + ".convf1_.bias"].to(dev).float().contiguous()
        self.init_mhs = sd[e + ".init_hidden_state"].to(dev).float().reshape(48).contiguous() if arch == "mof" else None
        self.w_v = _pad_k(sd["update_block.aggregator.to_v.weight"].to(dev).flatten(1).to(dt))
        self.gamma = float(sd["update_block.aggregator.gamma"].reshape(-1)[0])
        self.w_qk = _pad_k(sd["att.to_qk.weight"].to(dev).flatten(1).to(dt))
        # mask head (stack.py:120-124): conv3x3 128->256 as im2col GEMM with tap-major K, 1x1 256->1152, x100
        m0 = sd["update_block.mask.0.weight"].to(dev)                       # [256, 128, 3, 3]
        self.w_m0 = _pad_k(m0.permute(0, 2, 3, 1).reshape(256, 9 * 128).to(dt))
        self.b_m0 = sd["update_block.mask.0.bias"].to(dev).float().contiguous()
        self.w_m2 = _pad_k(sd["update_block.mask.2.weight"].to(dev).flatten(1).to(dt))
        self.mask_scale = 100.0 if arch == "mof" else 0.25          # stack.py:139 / sk2.py:91
        self.b_m2_x100 = (self.mask_scale * sd["update_block.mask.2.bias"].to(dev).float()).contiguous()
        enc_dt = encoder_dtype or torch.float32
        self.fnet = TwinsStage01(sd, "fnet.svt", dev, enc_dt) if "fnet.svt.patch_embeds.0.proj.weight" in sd else None
        self.cnet = TwinsStage01(sd, "cnet.svt", dev, enc_dt) if "cnet.svt.patch_embeds.0.proj.weight" in sd else None
        self._ws = {}                    # (Fc, h, w) -> workspace; a small LRU (tile mode alternates between 2-3 shapes)
        self.max_workspaces = int(os.environ.get("VFML_MAX_WORKSPACES", "3"))
        self.use_graph = use_graph
        self.vol_dtype = None            # untiled tiers only: storage type of the volume (default: the tier's dtype)
        self.lookup_math_f32 = False     # fp16x2 interpolation: 2.43e-2 px vs 2.37e-2 px (fp32) end to end, 8 % faster
        self.graph_launches = 0          # kernels executed through CUDA-graph replays (not seen by vfml_launch_count)
        # re-use of correlation pyramids between consecutive sliding windows (refine(frame_keys=...)); VFML_PYRAMID_CACHE=0 disables
        self.pyramid_cache = os.environ.get("VFML_PYRAMID_CACHE", "1") != "0"
        self.pyramid_slides = 0          # windows that re-used their predecessor's pyramids
        # ring order: a sliding window leaves the per-frame state of the shared centre frames (pyramids, attention matrices) in
        # the slots it is in and only the new frame's slot is computed; slot f holds the window's centre frame (f - rot) mod Fc.
        # VFML_PYRAMID_RING=0: the pyramids are moved down one slot instead (device copies) and the attention is recomputed.
        self.pyramid_ring = os.environ.get("VFML_PYRAMID_RING", "1") != "0"
        # build the pyramids on a side stream beside the attention: measured, no difference at C1 (206.2 +- 0.2 ms per 8 windows
        # either way: the prefetched encoders already fill that phase), so off
        self.pyramid_side_stream = os.environ.get("VFML_PYRAMID_SIDE", "0") == "1"
        self.encode_per_frame = os.environ.get("VFML_ENCODE_PER_FRAME", "1") != "0"      # see encode_new_frames
        self.use_side_stream = os.environ.get("VFML_NO_SIDE_STREAM") is None

    # ------------------------------------------------------------------ buffers
    def _rows_buf(self, n, c, dtype=None):
        pitch = _pad_act(c)
        return torch.zeros((n, pitch), dtype=dtype or self.dt, device=self.dev)[:, :c]

    def _workspace(self, Fc, h, w):
        """Buffers (and, once captured, the CUDA graph) of one window shape.  Kept in an LRU of `max_workspaces` shapes:
        a ragged tile grid (4K: 1280x1280 and 1280x880 tiles) or a change of sequence length switches shape without
        dropping the other shape's multi-GB workspace and re-capturing its ~1300-launch graph."""
        key = (Fc, h, w)
        ws = self._ws.pop(key, None)
        if ws is not None:
            self._ws[key] = ws            # most recently used last
            return ws
        while len(self._ws) >= max(1, self.max_workspaces):
            old = self._ws.pop(next(iter(self._ws)))
            old.clear()
            del old
        if True:
            Np = Fc * h * w
            ws = {
                "corr2": self._rows_buf(2 * Np, 324),
                "cor_in": self._rows_buf(Np, 256),
                "x400": self._rows_buf(Np, 400),      # MOF: [cor 192 | flo 64 | warped next 48 | warped prev 48 | hidden 48]
                "x256": self._rows_buf(Np, 256) if self.arch == "bof" else None,     # BOF: [cor 192 | flo 64]
                "f128": self._rows_buf(Np, 128),
                "x512": self._rows_buf(Np, 512),
                "vT": torch.zeros((Fc, 128, _pad8(h * w)), dtype=self.dt, device=self.dev),
                "delta": self._rows_buf(Np, 4, torch.float32),
                "sa": self._rows_buf(2 * Np, 512),
                "sb": self._rows_buf(2 * Np, 512),
                "hid": self._rows_buf(2 * Np, 768),
                "hid2": self._rows_buf(Np, 192), "sa2": self._rows_buf(Np, 128), "sb2": self._rows_buf(Np, 128),   # side stream (convf2)
                "qk": self._rows_buf(Np, 256),
                "attn": torch.zeros((Fc, h * w, _pad8(h * w)), dtype=self.dt, device=self.dev),
                "logits": torch.empty((h * w, h * w), dtype=torch.float32, device=self.dev),
                "mask": self._rows_buf(Np, 576, torch.float32),
            }
            if self.tiled and (self.dev.type == "cuda" or hasattr(self.ops, "tiled_pyramid_alloc")):
                # pyramids and the volume workspace live here, so the eager warm-up and the captured graph use the same
                # memory (not one copy in the caching allocator and another in the graph's private pool)
                ws["fpyr"] = self.ops.tiled_pyramid_alloc(Fc, h, w, self.dev)
                ws["bpyr"] = self.ops.tiled_pyramid_alloc(Fc, h, w, self.dev)
                ws["vol_ws"] = self.ops.tiled_volume_workspace(Fc, 256, h, w, self.dev)
            # split-K partial sums of the aggregate, sized once for the worst split of this shape: graphs captured earlier
            # keep the raw pointer, so the buffer may never be replaced afterwards
            P = h * w
            s_max = max([self._splits(P, k * P) for k in range(1, 4)] + [1])
            for f in range(Fc):
                ws[f"partial{f}"] = torch.empty(s_max * P * 128, dtype=torch.float32, device=self.dev)
            self._ws[key] = ws
        return ws

    @staticmethod
    def _view(buf, n, c):
        """[n, c] rows view carved from the start of a scratch buffer, pitch = _pad_act(c)."""
        base = buf._base if buf._base is not None else buf
        flat = base.reshape(-1)
        pitch = _pad_act(c)
        if n * pitch > flat.numel():
            raise RuntimeError("scratch buffer too small")
        return flat[: n * pitch].view(n, pitch)[:, :c]

    # ------------------------------------------------------------------ side stream (fork / join inside the captured graph)
    def _fork_side(self, k=0):
        """Following launches go to side stream k (ordered after everything issued on the main stream so far)."""
        self._side_ctx = None
        if self.dev.type != "cuda" or not getattr(self, "use_side_stream", True):
            return
        sides = self.__dict__.setdefault("_sides", {})
        if k not in sides:
            sides[k] = torch.cuda.Stream(device=self.dev)
        sides[k].wait_stream(torch.cuda.current_stream(self.dev))
        self._side_ctx = torch.cuda.stream(sides[k])
        self._side_ctx.__enter__()
        self.__dict__.setdefault("_forked", set()).add(k)

    def _back_to_main(self):
        if self._side_ctx is not None:
            self._side_ctx.__exit__(None, None, None)
            self._side_ctx = None

    def _join_side(self, k=0):
        """The main stream waits for side stream k."""
        if k in self.__dict__.get("_forked", ()):
            torch.cuda.current_stream(self.dev).wait_stream(self._sides[k])
            self._forked.discard(k)

    # ------------------------------------------------------------------ aggregate
    @staticmethod
    def _splits(M, K):
        """Split-K factor of the aggregate / read-out GEMM `attn [M, K] @ v [K, 128]`: its M / 128 row tiles (55 at
        config C1) leave most of the 148 SMs idle while each busy SM streams 1.8 MB of attention rows, so K is cut into
        S equal multiples of 64 until most SMs have a CTA (measured at C1: S = 2, 5, 10 -> 29, 30, 35 us against 40 us unsplit;
        the partial sums cost S x 3.6 MB per frame).  1 = not worth it / not divisible."""
        m_tiles = (M + 127) // 128
        if K % 64 or m_tiles >= _sm_count():
            return 1
        nkb = K // 64
        for S in range(2, nkb + 1):
            if nkb % S == 0 and K // S >= 512 and m_tiles * S >= 100 * _sm_count() // 148:
                return S
        return 1

    def _aggregate(self, ws, attn, vT, res, out, slot=0):
        """out = res + gamma * attn @ vT^T   (gma.py:113-115; memory read-out + inference_core_skflow.py:43).
        slot: which partial-sum buffer to use (frames that run concurrently on different streams need their own)."""
        M, K = attn.shape
        S = self._splits(M, K) if hasattr(self.ops, "gemm_splitk") else 1
        if S == 1:
            return self.ops.gemm(attn, vT, out, res=res, alpha=self.gamma)
        key = f"partial{slot}"
        part = ws.get(key)
        if part is None or part.numel() < S * M * out.shape[1]:
            if ws.get("graph") is not None or ws.get("graphs"):
                raise RuntimeError("split-K partial buffer too small after graph capture (captured graphs hold its address)")
            part = ws[key] = torch.empty(S * M * out.shape[1], dtype=torch.float32, device=self.dev)
        return self.ops.gemm_splitk(attn, vT, out, S, part, res=res, alpha=self.gamma)

    # ------------------------------------------------------------------ PCBlock
    def _pcblock(self, x, W, ws, n_img, h, w, outs, scratch=("hid", "sa", "sb")):
        """x: rows view [M, Cin].  outs: list of (weight_row_slice, row_slice, out_view, act) for the
        last 1x1 conv, so that the reference's cat/split around the block costs nothing.
        scratch: names of the three workspace buffers the block may use (a second set for the side stream)."""
        o = self.ops
        M = x.shape[0]
        ws = {"hid": ws[scratch[0]], "sa": ws[scratch[1]], "sb": ws[scratch[2]]}
        t1 = self._view(ws["hid"], M, W.hid)
        o.gemm(x, W.w1a, t1, bias=W.b1a, act="gelu", weights=True)                                         # ffn1.0 + GELU
        x2 = self._view(ws["sa"], M, W.cin)
        o.gemm(t1, W.w1b, x2, bias=W.b1b, res=x, act="gelu", dw=(W.dw1_w, W.dw1_b), weights=True)            # gelu(x+ffn1) ; k=1 dw stage
        x3 = self._view(ws["sb"], M, W.cin)
        if W.dwk_packed is not None:                                                           # gelu(x + dw_k(x))
            o.dwconv_tc_res_gelu(x2, x3, n_img, h, w, W.dwk_packed, W.dwk_b, W.k)
        else:
            o.dwconv_res_gelu(x2, x3, n_img, h, w, W.dwk_w, W.dwk_b, W.k)
        x4 = self._view(ws["sa"], M, W.cin)
        o.gemm(x3, W.wpw, x4, bias=W.bpw, res=x3, act="gelu", weights=True)                                  # gelu(x + pw(x))
        t2 = self._view(ws["hid"], M, W.hid)
        o.gemm(x4, W.w2a, t2, bias=W.b2a, act="gelu", weights=True)                                          # ffn2.0 + GELU
        for wsl, rsl, out, act in outs:                                                         # ffn2.2
            o.gemm(t2[rsl], W.w2b[wsl], out, bias=W.b2b[wsl].contiguous(), act=act, weights=True)

    # ------------------------------------------------------------------ pieces
    def _guard(self):
        """Every launch below goes to the current stream of the CURRENT device: make that the engine's device (the
        reference wrappers accept any `device` without the caller calling torch.cuda.set_device first)."""
        import contextlib
        return torch.cuda.device(self.dev) if self.dev.type == "cuda" else contextlib.nullcontext()

    @torch.no_grad()
    def encode(self, images):
        """images [1, T, 3, H, W] (raw caller values) -> fmaps [T,256,h,w], net, inp [T-2,128,h,w].
        network.py:144-170 (the 2*(x/255)-1 normalisation stays inside)."""
        with self._guard():
            return self._encode(images)

    def _encode(self, images):
        B, T, _, H, W = images.shape
        x = 2 * (images.float() / 255.0) - 1.0
        self._fork_side()                                   # the context encoder runs beside the feature encoder
        cnet = self.cnet(x[:, 1:T - 1].reshape(B * (T - 2), 3, H, W))
        self._back_to_main()
        fmaps = self.fnet(x.reshape(B * T, 3, H, W))
        self._join_side()
        if cnet.is_cuda:
            cnet.record_stream(torch.cuda.current_stream(cnet.device))
        net, inp = torch.split(cnet, [128, 128], dim=1)
        return fmaps, torch.tanh(net).contiguous(), torch.relu(inp).contiguous()

    @torch.no_grad()
    def encode_frames(self, frames, context=False):
        """Per-frame halves of `encode` for the feature cache (fnet and cnet are per-image functions, network.py:150,161):
        frames [n, 3, H, W] (raw caller values, already padded) -> fmaps [n, 256, h, w] (context=False) or the context
        encoder's (tanh(net), relu(inp)) [n, 128, h, w] each (context=True)."""
        fm, ctx = self.encode_new_frames(None if context else frames, frames if context else None)
        return ctx if context else fm

    def _encode_new_body(self, xf, xc):
        fm = ctx = None
        if xc is not None:
            if xf is not None:
                self._fork_side()                           # the context encoder runs beside the feature encoder
            net, inp = torch.split(self.cnet(2 * (xc.float() / 255.0) - 1.0), [128, 128], dim=1)
            ctx = (torch.tanh(net).float().contiguous(), torch.relu(inp).float().contiguous())
            if xf is not None:
                self._back_to_main()
        if xf is not None:
            fm = self.fnet(2 * (xf.float() / 255.0) - 1.0).float()
            if xc is not None:
                self._join_side()
                for t in ctx:
                    if t.is_cuda:
                        t.record_stream(torch.cuda.current_stream(t.device))
        return fm, ctx

    @torch.no_grad()
    def encode_new_frames(self, xf, xc):
        """Encoders on the frames a window adds to the cache: xf [nf, 3, H, W] -> fmaps [nf, 256, h, w], xc [nc, 3, H, W] ->
        (tanh(net), relu(inp)) [nc, 128, h, w] each; either may be None.  The two encoders run on two streams, and the
        steady-state shapes (one new frame each: ~60 small library launches per encoder) are replayed from a CUDA graph."""
        with self._guard():
            nf, nc = (0 if xf is None else xf.shape[0]), (0 if xc is None else xc.shape[0])
            if self.use_graph and self.dev.type == "cuda" and max(nf, nc) > 1 and self.encode_per_frame:
                # a cold window (T + T-2 frames) goes through the same one-frame graphs as the steady state, frame by frame: no
                # large-batch library workspaces to allocate (sporadic ~0.4 s stalls of the first window of a clip were traced to
                # that path), the same kernels -- hence bit-identical features -- whether a frame is encoded alone or in a window
                fms, nets, inps = [], [], []
                for i in range(max(nf, nc)):
                    fm_i, ctx_i = self.encode_new_frames(xf[i:i + 1] if i < nf else None, xc[i:i + 1] if i < nc else None)
                    if fm_i is not None:
                        fms.append(fm_i)
                    if ctx_i is not None:
                        nets.append(ctx_i[0]); inps.append(ctx_i[1])
                return (torch.cat(fms) if fms else None), ((torch.cat(nets), torch.cat(inps)) if nets else None)
            if not (self.use_graph and self.dev.type == "cuda" and max(nf, nc) <= 1):
                return self._encode_new_body(xf, xc)
            key = (None if xf is None else tuple(xf.shape), None if xc is None else tuple(xc.shape))
            graphs = self.__dict__.setdefault("_enc_graphs", {})
            g = graphs.pop(key, None)
            if g is None:
                while len(graphs) >= 12:        # 3 encoder combinations (fnet + cnet, fnet only, cnet only) x the tile shapes of a clip
                    graphs.pop(next(iter(graphs)))
                static = [None if t is None else t.float().clone() for t in (xf, xc)]
                self._encode_new_body(*static)              # eager warm-up (library handles, autotuning) before the capture
                torch.cuda.synchronize(self.dev)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    out = self._encode_new_body(*static)
                g = {"graph": graph, "static": static, "out": out}
            graphs[key] = g
            for s_, t in zip(g["static"], (xf, xc)):
                if t is not None:
                    s_.copy_(t)
            g["graph"].replay()
            fm, ctx = g["out"]       # static outputs: the cache keeps copies
            return (None if fm is None else fm.clone()), (None if ctx is None else (ctx[0].clone(), ctx[1].clone()))

    def build_pyramids(self, fmaps, ws=None, slide=False, rot=0):
        """Forward (frames 1..T-2 vs 2..T-1) and backward (vs 0..T-3) pyramids, network.py:157-158.
        slide=True (tiled pyramids kept in the workspace): the previous window was this one shifted by one frame, so its
        pyramids of centre frames 1.. ARE this window's pyramids of centre frames 0.. (a correlation pyramid depends on its
        two frames only): they move down one slot (device copies, 4x cheaper than the volume GEMM + pooling) and only the
        last centre frame's pair is computed."""
        T = fmaps.shape[0]
        if slide:
            Fc = T - 2
            fp, bp = ws["fpyr"], ws["bpyr"]
            P = fp.h * fp.w
            if self.pyramid_ring:
                sn = (Fc - 1 + rot) % Fc             # the slot of the frame that left the window
            else:
                sn = Fc - 1
                for pyr in (fp, bp):
                    for lv in pyr.levels:
                        for j in range(Fc - 1):      # ascending: slot j is free once it has been copied to j - 1
                            lv[j * P:(j + 1) * P].copy_(lv[(j + 1) * P:(j + 2) * P])
            last = lambda pyr: self.ops.TiledPyramid([lv[sn * P:(sn + 1) * P] for lv in pyr.levels], 1, pyr.h, pyr.w)
            c1 = fmaps[T - 2:T - 1].contiguous()
            self.ops.corr_volume_pyramid_tiled(c1, fmaps[T - 1:T].contiguous(), ws.get("vol_ws"), out=last(fp))
            self.ops.corr_volume_pyramid_tiled(c1, fmaps[T - 3:T - 2].contiguous(), ws.get("vol_ws"), out=last(bp))
            return fp, bp
        c = fmaps[1:T - 1].contiguous()
        if self.tiled:
            ws = ws or {}
            return (self.ops.corr_volume_pyramid_tiled(c, fmaps[2:T].contiguous(), ws.get("vol_ws"), out=ws.get("fpyr")),
                    self.ops.corr_volume_pyramid_tiled(c, fmaps[0:T - 2].contiguous(), ws.get("vol_ws"), out=ws.get("bpyr")))
        vdt = self.vol_dtype or self.dt
        fw = self.ops.corr_volume_pyramid(c, fmaps[2:T].contiguous(), CORR_LEVELS, vdt)
        bw = self.ops.corr_volume_pyramid(c, fmaps[0:T - 2].contiguous(), CORR_LEVELS, vdt)
        return fw, bw

    def attention(self, ws, Fc, P, slots=None):
        """Attention.forward (gma.py:54-76): softmax(q k^T / sqrt(128)) per centre frame (of the given slots: a sliding window
        keeps the matrices of the frames it shares with its predecessor).  One GEMM shape per frame whatever the number of
        slots, so a matrix does not depend on which window computed it."""
        o = self.ops
        inp = ws["x512"][:, 128:256]
        for f in (range(Fc) if slots is None else slots):
            o.gemm(inp[f * P:(f + 1) * P], self.w_qk, ws["qk"][f * P:(f + 1) * P])
            q = ws["qk"][f * P:(f + 1) * P, 0:128]
            k = ws["qk"][f * P:(f + 1) * P, 128:256]
            o.gemm(q, k, ws["logits"], alpha=128 ** -0.5)
            o.softmax_rows(ws["logits"], ws["attn"][f][:, :P])

    def update_step(self, ws, fpyr, bpyr, fc, bc, Fc, h, w, rot=0):
        """One iteration of network.py:180-197 (lookups + SKUpdateBlock6..AllDecoder2, stack.py:127-140)."""
        o = self.ops
        P = h * w
        Np = Fc * P
        corr2, x400, x512 = ws["corr2"], ws["x400"], ws["x512"]
        mf = x512[:, 256:384]
        full = slice(None)
        xin = ws["x256"] if self.arch == "bof" else x400
        # the flow branch (glue + convf2, stack.py:81-87,94-95) is independent of the correlation branch: it runs on a side
        # stream and fills the partial last waves of the big convc1 / convc2 kernels
        self._fork_side()
        o.motion_glue(fc, bc, x400, ws["f128"], mf, self.convf1_w, self.convf1_b, Fc, h, w, rot=rot)
        self._pcblock(ws["f128"], self.convf2, ws, Fc, h, w, [(full, full, xin[:, 192:256], None)], scratch=("hid2", "sa2", "sb2"))
        self._back_to_main()
        if self.tiled:
            o.corr_lookup_tiled(fpyr, bpyr, fc, bc, corr2[:Np], corr2[Np:], math_f32=self.lookup_math_f32)
        else:
            o.corr_lookup_pair(fpyr, bpyr, fc, bc, corr2[:Np], corr2[Np:])
        # convc1 on forward and backward correlation (shared weights), gelu(cat) -> cor_in   stack.py:89-92
        self._pcblock(corr2, self.convc1, ws, 2 * Fc, h, w,
                      [(full, slice(0, Np), ws["cor_in"][:, 0:128], "gelu"),
                       (full, slice(Np, 2 * Np), ws["cor_in"][:, 128:256], "gelu")])
        xin = ws["x256"] if self.arch == "bof" else x400     # BOF has no hidden state: the glue's warps land in a scratch x400
        self._pcblock(ws["cor_in"], self.convc2, ws, Fc, h, w, [(full, full, xin[:, 0:192], None)])
        self._join_side()
        if self.arch == "bof":
            self._pcblock(xin, self.conv, ws, Fc, h, w, [(full, full, x512[:, 256:380], None)])     # sk2.py:55-56
        else:
            # conv: 400 -> 172 = 124 motion features + 48 new hidden state                     stack.py:97-102
            self._pcblock(x400, self.conv, ws, Fc, h, w,
                          [(slice(0, 124), full, x512[:, 256:380], None),
                           (slice(124, 172), full, x400[:, 352:400], None)])
        # Aggregate (gma.py:102-115): mfg = mf + gamma * attn @ to_v(mf)
        # the centre frames are independent and each one's kernels fill well under 148 SMs (55 / 110 CTAs):
        # frame 0 stays on the main stream, the others run beside it
        for f in list(range(1, Fc)) + [0]:          # side frames first: a fork orders the side stream after the main stream's past
            rows = slice(f * P, (f + 1) * P)
            vT = ws["vT"][f][:, :P]
            if f:
                self._fork_side(f)
            o.gemm(self.w_v, mf[rows], vT)                                   # to_v, transposed: [128, P]
            self._aggregate(ws, ws["attn"][f][:, :P], vT, mf[rows], x512[rows, 384:512], slot=f)
            if f:
                self._back_to_main()
        for f in range(1, Fc):
            self._join_side(f)
        self._pcblock(x512, self.gru, ws, Fc, h, w, [(full, full, x512[:, 0:128], None)])
        self._pcblock(x512[:, 0:128], self.fhead, ws, Fc, h, w, [(full, full, ws["delta"], None)])
        o.coords_add(fc, bc, ws["delta"])

    def mask_rows(self, ws, frames, h, w, chan):
        """mask head on the rows of `frames` (a slice of centre frames): 100 * mask(net)[:, chan]. stack.py:139"""
        o = self.ops
        P = h * w
        n_img = frames.stop - frames.start
        rows = slice(frames.start * P, frames.stop * P)
        net = ws["x512"][rows, 0:128]
        col = self._view(ws["hid"], n_img * P, 9 * 128)
        o.im2col3x3(net, col, n_img, h, w)
        m256 = self._view(ws["sa"], n_img * P, 256)
        o.gemm(col, self.w_m0, m256, bias=self.b_m0, act="relu")
        nch = chan.stop - chan.start
        mask = ws["mask"][: n_img * P, :nch]       # fp32 in both tiers: logits of the 9-tap softmax
        o.gemm(m256, self.w_m2[chan], mask, bias=self.b_m2_x100[chan].contiguous(), alpha=self.mask_scale)
        return mask

    def _mask_chan(self, direction):
        """Mask-head channels of the forward (0) / backward (1) flow: MOF [fwd | bwd] (network.py:199),
        BOF [2->1 | 2->3] = [bwd | fwd] (BOFNet/network.py:130)."""
        d = direction if self.arch == "mof" else 1 - direction
        return slice(576 * d, 576 * (d + 1))

    def _iterations_body(self, ws, fmaps, net, inp, Fc, h, w, iters=None, slide=False, rot=0):
        """pyramids, attention, `iters` update steps.  Leaves net in x512[:, 0:128]; returns the coords.  Every per-frame buffer
        is in ring order: slot f = the window's centre frame (f - rot) mod Fc (rot = 0 unless the window slid, pyramid_ring)."""
        o = self.ops
        P = h * w
        # the pyramids (store-bound volume kernel) are independent of the context rows + attention (GEMM + softmax): side stream
        side = self.pyramid_side_stream and self.tiled and ws.get("fpyr") is not None
        if side:
            self._fork_side()
        fpyr, bpyr = self.build_pyramids(fmaps, ws, slide=slide, rot=rot)
        if side:
            self._back_to_main()
        x512 = ws["x512"]
        if rot:
            k = Fc - rot                  # centre frames k.. live in slots 0 .. rot-1, frames 0 .. k-1 in slots rot ..
            for src, cols in ((net, slice(0, 128)), (inp, slice(128, 256))):
                o.nchw_to_rows(src[k:], x512[:rot * P, cols])
                o.nchw_to_rows(src[:k], x512[rot * P:, cols])
        else:
            o.nchw_to_rows(net, x512[:, 0:128])
            o.nchw_to_rows(inp, x512[:, 128:256])
        if self.arch == "mof":
            o.rows_fill(ws["x400"][:, 352:400], self.init_mhs)      # init_hidden_state, stack.py:72-73
        self.attention(ws, Fc, P, [(Fc - 1 + rot) % Fc] if (slide and self.pyramid_ring) else None)
        if side:
            self._join_side()
        fc = o.coords_grid(Fc, h, w, self.dev)
        bc = o.coords_grid(Fc, h, w, self.dev)
        for _ in range(self.iters if iters is None else iters):
            self.update_step(ws, fpyr, bpyr, fc, bc, Fc, h, w, rot=rot)
        return fc, bc

    @staticmethod
    def _pair_keys(frame_keys, Fc):
        """identity of every pyramid of a window: (centre frame, partner frame) per slot, forward then backward"""
        if frame_keys is None or any(k is None for k in frame_keys):
            return None
        k = list(frame_keys)
        return tuple((k[j + 1], k[j + 2]) for j in range(Fc)) + tuple((k[j + 1], k[j]) for j in range(Fc))

    def _run_iterations(self, ws, fmaps, net, inp, Fc, h, w, frame_keys=None):
        """The ~1300 launches of a window are captured once per shape into a CUDA graph and replayed
        (inputs are copied into static buffers; the graph owns the pyramids and the coords).  A second graph per shape
        serves windows that follow their predecessor by one frame (frame_keys: hashable identities of the T frames, the
        same keys as the per-frame feature cache): it re-uses the pyramids of the shared centre frames (build_pyramids)."""
        fmaps, net, inp = fmaps.float().contiguous(), net.float().contiguous(), inp.float().contiguous()
        pairs = self._pair_keys(frame_keys, Fc) if (self.tiled and self.pyramid_cache and Fc > 1) else None
        prev = ws.get("pyr_keys")
        slide = (pairs is not None and prev is not None
                 and all(pairs[j] == prev[j + 1] and pairs[Fc + j] == prev[Fc + j + 1] for j in range(Fc - 1)))
        ws["pyr_keys"] = None                      # (stays None if anything below raises: the slots are then in an unknown state)
        rot = (ws.get("rot", 0) + 1) % Fc if (slide and self.pyramid_ring) else 0
        if not (self.use_graph and fmaps.is_cuda):
            out = self._iterations_body(ws, fmaps, net, inp, Fc, h, w, slide=slide, rot=rot)
            ws["pyr_keys"], ws["rot"] = pairs, rot
            self.pyramid_slides += int(slide)
            return out, rot
        graphs = ws.setdefault("graphs", {})
        g = graphs.get((slide, rot))
        if g is None or g["iters"] != self.iters:
            static = [torch.empty_like(t) for t in (fmaps, net, inp)]
            for s_, t in zip(static, (fmaps, net, inp)):
                s_.copy_(t)
            if not graphs:
                with _ops.gemm_autotuning():                            # each layer shape picks its GEMM kernel by measurement, once
                    self._iterations_body(ws, *static, Fc, h, w, iters=1)   # eager warm-up (one iteration reaches every kernel): module
            torch.cuda.synchronize(self.dev)                        # loading and one-time attributes cannot happen under capture
            # (a first slide-graph capture leaves the previous window's pyramids in place: capturing records, it does not run)
            graph = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with torch.cuda.graph(graph):
                coords = self._iterations_body(ws, *static, Fc, h, w, slide=slide, rot=rot)
            g = graphs[(slide, rot)] = {"graph": graph, "static": static, "coords": coords, "iters": self.iters,
                                        "launches": _lib.launch_count() - n0}
        for s_, t in zip(g["static"], (fmaps, net, inp)):
            s_.copy_(t)
        g["graph"].replay()
        ws["pyr_keys"], ws["rot"] = pairs, rot
        self.pyramid_slides += int(slide)
        self.graph_launches += g["launches"]
        return g["coords"], rot

    # ------------------------------------------------------------------ forward
    @torch.no_grad()
    def refine(self, fmaps, net, inp, select=None, crop=None, out=None, frame_keys=None):
        """fmaps [T,256,h,w], net/inp [T-2,128,h,w] (fp32, CUDA).
        select=None  -> all flows [1, 2(T-2), 2, 8h, 8w] (forward flows then backward), network.py:210-215
        select=index -> only that flow [2, out_h, out_w]; crop=(top, left, out_h, out_w) fuses the unpad."""
        with self._guard():
            return self._refine(fmaps, net, inp, select, crop, out, frame_keys)

    def _refine(self, fmaps, net, inp, select, crop, out, frame_keys=None):
        o = self.ops
        T, _, h, w = fmaps.shape
        Fc, P = T - 2, h * w
        ws = self._workspace(Fc, h, w)
        (fc, bc), rot = self._run_iterations(ws, fmaps, net, inp, Fc, h, w, frame_keys)
        if select is None:
            flows = torch.empty((1, 2 * Fc, 2, 8 * h, 8 * w), dtype=torch.float32, device=self.dev)
            for d, coords in enumerate((fc, bc)):
                if rot == 0:
                    mk = self.mask_rows(ws, slice(0, Fc), h, w, self._mask_chan(d))
                    o.convex_upsample8(coords, mk, out=flows[0, d * Fc:(d + 1) * Fc], flow_is_coords=True)
                    continue
                for j in range(Fc):                   # ring order: the window's centre frame j is in slot (j + rot) mod Fc
                    sl = (j + rot) % Fc
                    mk = self.mask_rows(ws, slice(sl, sl + 1), h, w, self._mask_chan(d))
                    o.convex_upsample8(coords[sl:sl + 1].contiguous(), mk, out=flows[0, d * Fc + j:d * Fc + j + 1], flow_is_coords=True)
            return flows
        direction, frame = divmod(select, Fc)
        frame = (frame + rot) % Fc                    # its slot
        coords = (fc, bc)[direction][frame:frame + 1]
        mk = self.mask_rows(ws, slice(frame, frame + 1), h, w, self._mask_chan(direction))
        if out is None:
            oh, ow = (crop[2], crop[3]) if crop else (8 * h, 8 * w)
            out = torch.empty((2, oh, ow), dtype=torch.float32, device=self.dev)
        o.convex_upsample8(coords.contiguous(), mk, out=out[None], crop=crop, flow_is_coords=True)
        return out

    @torch.no_grad()
    def forward(self, images, select=None, crop=None):
        fmaps, net, inp = self.encode(images)
        return self.refine(fmaps, net, inp, select=select, crop=crop)
