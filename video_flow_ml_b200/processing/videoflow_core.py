"""VideoFlowCore on libvfml kernels.  Mirrors processing/videoflow_core.py:33-242 of the reference:
same constructor, methods, checkpoint naming, return values and exceptions."""
import collections
import os

import torch

from ..mof import MOFEngine

_PRECISIONS = {"fp32": torch.float32, "fp16": torch.float16}


def pad_amounts(ht, wd, mode="sintel"):
    """InputPadder (VideoFlow/core/utils/utils.py:9-18): (left, right, top, bottom) to a multiple of 8."""
    ph = (((ht // 8) + 1) * 8 - ht) % 8
    pw = (((wd // 8) + 1) * 8 - wd) % 8
    if mode == "sintel":
        return pw // 2, pw - pw // 2, ph // 2, ph - ph // 2
    return pw // 2, pw - pw // 2, 0, ph


class _Config(dict):
    """Minimal attribute + item config (the reference uses a yacs CfgNode)."""
    __getattr__ = dict.get

    def __setattr__(self, k, v):
        self[k] = v


class _FeatureArena:
    """Storage of the per-frame feature cache for ONE feature-map shape: `cap` slots of (fmap, net, inp), each of the three
    allocated once.  A growing cache would otherwise take ~14 MB per new frame from the caching allocator, i.e. from
    cudaMalloc until the cache is full (the first 96 frames of a clip): measured on the B200 VMs, those calls stall the
    launching thread for 70 - 600 ms every few windows (profiles/r2_bench_host_stalls.txt).  Slots of evicted frames are
    reused in place; writes and reads of a slot are ordered by the stream they are issued on."""

    def __init__(self, cap):
        self.cap = cap
        self.free = list(range(cap - 1, -1, -1))
        self.fmap = self.net = self.inp = None

    def put_fmap(self, slot, fm):
        if self.fmap is None:
            self.fmap = torch.empty((self.cap,) + tuple(fm.shape), dtype=fm.dtype, device=fm.device)
        self.fmap[slot].copy_(fm)
        return self.fmap[slot]

    def put_ctx(self, slot, net, inp):
        if self.net is None:
            self.net = torch.empty((self.cap,) + tuple(net.shape), dtype=net.dtype, device=net.device)
            self.inp = torch.empty((self.cap,) + tuple(inp.shape), dtype=inp.dtype, device=inp.device)
        self.net[slot].copy_(net)
        self.inp[slot].copy_(inp)
        return self.net[slot], self.inp[slot]


class VideoFlowCore:
    def __init__(self, device, fast_mode=False, dataset='sintel', architecture='mof', variant='standard',
                 precision=None):
        self.device = device
        self.fast_mode = fast_mode
        self.dataset = dataset
        self.architecture = architecture.lower()
        self.variant = variant
        self.precision = precision or os.environ.get("VFML_PRECISION", "fp32")
        if self.precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}, got {self.precision}")
        self.model = None
        self.cfg = None
        # per-frame feature cache: fnet / cnet are per-image functions (network.py:150,161), and sliding
        # windows share T-1 of their T frames, so a steady-state window encodes ONE new frame instead of T + (T-2)
        self.feature_cache_frames = int(os.environ.get("VFML_FEATURE_CACHE", "96"))
        self._feat = collections.OrderedDict()        # key -> {"fmap": [256,h,w], "ctx": (net, inp) or None}
        # CUDA: the cached tensors live in per-shape arenas (no allocation per new frame); an arena holds at most
        # feature_cache_frames slots and VFML_FEATURE_CACHE_MB (default 4096) megabytes, never fewer than two windows
        self.feature_cache_bytes = int(os.environ.get("VFML_FEATURE_CACHE_MB", "4096")) << 20
        self._arenas = {}
        self.cache_stats = {"fnet_runs": 0, "cnet_runs": 0, "fnet_hits": 0, "cnet_hits": 0}
        # encoder outputs of the NEXT window computed on a side stream while the current one iterates (VFML_PREFETCH=0 or
        # core.prefetch = False turns it off): at C1 it hides 0.2 ms of the 1.0 ms encoder pass in the device-resident loop and
        # 0.4 ms per call from host frames (profiles/r2_ring_fused_ab.txt).  It was opt-in while the feature cache still allocated
        # per frame: the cudaMalloc stalls of that version hit the prefetching runs hardest and were mistaken for its cost.
        self.prefetch = os.environ.get("VFML_PREFETCH", "1") == "1"
        self._pf_stream = None
        self._pf_event = None

    # ------------------------------------------------------------------ loading
    def _model_path(self):
        arch_upper = self.architecture.upper()
        if self.variant == 'noise' and self.dataset == 'things':
            return f"VideoFlow_ckpt/{arch_upper}_{self.dataset}_288960noise.pth"
        return f"VideoFlow_ckpt/{arch_upper}_{self.dataset}.pth"

    def _make_cfg(self, model_path):
        # VideoFlow/configs/multiframes_sintel_submission.py:16-37 (the values the network is built from)
        cfg = _Config(model=model_path, network="MOFNetStack", decoder_depth=32, corr_levels=4, corr_radius=4,
                      mixed_precision=self.precision == "fp16", feat_dim=256, down_ratio=8)
        if self.fast_mode:   # written on the top-level cfg only; the network ignores them
            cfg.decoder_depth, cfg.corr_levels, cfg.corr_radius = 6, 3, 3
        return cfg

    def load_model(self):
        model_path = self._model_path()
        if not os.path.exists(model_path):
            raise FileNotFoundError(f"VideoFlow model weights not found: {model_path}")
        checkpoint = torch.load(model_path, map_location="cpu")
        self.load_state_dict(checkpoint, model_path)
        print(f"[Model] VideoFlow model loaded successfully:")
        print(f"  Path: {model_path}")
        print(f"  Architecture: {self.architecture.upper()} ({self.architecture})")
        print(f"  Dataset: {self.dataset}")
        print(f"  Variant: {self.variant}")
        print(f"  Device: {self.device}")
        return model_path

    def load_state_dict(self, state_dict, model_path="<state_dict>", iters=32):
        """Build the engine from an in-memory checkpoint (upstream key names, optional 'module.' prefix)."""
        if self.architecture not in ('mof', 'bof'):
            raise ValueError(f"unknown architecture {self.architecture!r} (expected 'mof' or 'bof')")
        if not str(self.device).startswith("cuda"):
            raise RuntimeError("this build runs on CUDA devices only (no CPU fallback)")
        self.cfg = self._make_cfg(model_path)
        if self.architecture == 'bof' and iters == 32:
            iters = 12            # VideoFlow/configs/things.py:31 (BOFNet.decoder_depth)
        self.clear_feature_cache()
        self.model = MOFEngine(state_dict, self.device, _PRECISIONS[self.precision], iters=iters,
                               encoder_dtype=_PRECISIONS[self.precision] if self.precision == "fp16" else None,
                               arch=self.architecture)

    # ------------------------------------------------------------------ inference
    def _validate(self, t):
        if self.model is None:
            raise RuntimeError("VideoFlow model not loaded. Call load_model() first.")
        if not isinstance(t, torch.Tensor):
            raise ValueError("Input must be a torch.Tensor")
        td, md = str(t.device), str(self.device)
        if not ((td.startswith('cuda') and md == 'cuda') or (md.startswith('cuda') and td == 'cuda') or td == md):
            raise ValueError(f"Input tensor device ({t.device}) doesn't match model device ({self.device})")
        if len(t.shape) != 5:
            raise ValueError(f"Input tensor must have 5 dimensions [B,T,C,H,W], got {len(t.shape)}")
        if t.shape[0] != 1:
            raise ValueError(f"Batch size must be 1, got {t.shape[0]}")
        if t.shape[2] != 3:
            raise ValueError(f"Must have 3 color channels, got {t.shape[2]}")

    def compute_flow_from_tensor(self, frame_batch_tensor, out=None):
        """[1, T, 3, H, W] -> [2, H, W]: replicate-pad to x8 ('sintel' mode), forward, unpad, flow index T-2
        (the first backward flow).  Only that flow is upsampled; the unpad is fused into the
        upsampling kernel's cropped write."""
        self._validate(frame_batch_tensor)
        H, W = frame_batch_tensor.shape[-2:]
        l, r, t, b = pad_amounts(H, W)
        x = frame_batch_tensor
        if l or r or t or b:
            x = torch.nn.functional.pad(x, [l, r, t, b, 0, 0], mode='replicate')
        Tn = x.shape[1]
        with torch.no_grad():
            fmaps, net, inp = self.model.encode(x)
            return self.model.refine(fmaps, net, inp, select=Tn - 2, crop=(t, l, H, W), out=out)

    def compute_flow_from_frames(self, frames, keys, out=None, next_window=None):
        """Same result as compute_flow_from_tensor(stack(frames)[None]) for a window given frame by frame, with the
        encoders' outputs cached per frame.  frames: T tensors [3, H, W] on the device (values as the caller would stack
        them), or callables returning one (evaluated only for frames whose features are not cached, so a cached frame
        is not even uploaded); keys: T hashable frame identities (None = do not cache that frame).
        next_window: (frames, keys) of the window the caller will ask for next: its missing encoder outputs (one new frame
        per encoder in a sliding pass) are computed on a side stream WHILE this window iterates, and are cache hits then."""
        if self.model is None:
            raise RuntimeError("VideoFlow model not loaded. Call load_model() first.")
        Tn = len(frames)
        if Tn < 3 or len(keys) != Tn:
            raise ValueError("compute_flow_from_frames needs >= 3 frames and one key per frame")
        centre = range(1, Tn - 1)
        with torch.no_grad():
            ent = self._ensure_features(frames, keys)
            if next_window is not None and self.prefetch and self.feature_cache_frames > 0:
                self._prefetch(*next_window)
            H, W = ent[0]["hw"]
            l, r, tp, b = pad_amounts(H, W)
            fmaps = torch.stack([e["fmap"] for e in ent])
            net = torch.stack([ent[i]["ctx"][0] for i in centre])
            inp = torch.stack([ent[i]["ctx"][1] for i in centre])
            return self.model.refine(fmaps, net, inp, select=Tn - 2, crop=(tp, l, H, W), out=out, frame_keys=keys)

    def _prefetch(self, frames, keys):
        """_ensure_features for a future window on the prefetch stream (ordered after everything queued so far on the
        current stream; the current stream waits for it at its next _ensure_features)."""
        if len(frames) != len(keys) or len(frames) < 3 or any(k is None for k in keys) or not torch.cuda.is_available():
            return
        main = torch.cuda.current_stream(self.device)
        if self._pf_stream is None:
            self._pf_stream = torch.cuda.Stream(self.device)
        ev = torch.cuda.Event()
        ev.record(main)
        with torch.cuda.stream(self._pf_stream):
            self._pf_stream.wait_event(ev)
            self._ensure_features(frames, keys, prefetch=True)
            self._pf_event = torch.cuda.Event()
            self._pf_event.record(self._pf_stream)

    def _ensure_features(self, frames, keys, prefetch=False):
        """Cache entries ({"fmap", "ctx", "hw"}) of the window's frames, running the encoders on what is missing."""
        Tn = len(frames)
        centre = range(1, Tn - 1)
        cur = torch.cuda.current_stream(self.device) if (torch.cuda.is_available() and str(self.device).startswith("cuda")) else None
        if not prefetch and self._pf_event is not None:
            cur.wait_event(self._pf_event)               # features (and the encoders' static buffers) the prefetch stream is writing
            self._pf_event = None
        ent = [self._feat.get(k) if k is not None else None for k in keys]
        need_f = [i for i in range(Tn) if ent[i] is None]
        need_c = [i for i in centre if ent[i] is None or ent[i]["ctx"] is None]
        # repeated frames of a padded window (clip start / end) are encoded once
        first = {}
        for i in sorted(set(need_f) | set(need_c)):
            first.setdefault(keys[i] if keys[i] is not None else ("#", i), i)
        loaded = {}

        def tensor(i):
            j = first[keys[i] if keys[i] is not None else ("#", i)]
            if j not in loaded:
                t = frames[j]() if callable(frames[j]) else frames[j]
                H, W = t.shape[-2:]
                l, r, tp, b = pad_amounts(H, W)
                t = t[None]
                if l or r or tp or b:
                    t = torch.nn.functional.pad(t, [l, r, tp, b], mode='replicate')
                loaded[j] = (t, H, W)
            return loaded[j]
        uniq_f = sorted({first[keys[i] if keys[i] is not None else ("#", i)] for i in need_f})
        uniq_c = sorted({first[keys[i] if keys[i] is not None else ("#", i)] for i in need_c})
        new = {}
        same = uniq_f == uniq_c and len(uniq_f) > 0
        xf = torch.cat([tensor(j)[0] for j in uniq_f]) if uniq_f else None
        xc = xf if same else (torch.cat([tensor(j)[0] for j in uniq_c]) if uniq_c else None)
        fm, ctx = self.model.encode_new_frames(xf, xc) if (uniq_f or uniq_c) else (None, None)      # the two encoders side by side
        pooled = cur is not None and self.feature_cache_frames > 0
        keep = [e for e in ent if e is not None]
        for n, j in enumerate(uniq_f):
            new[j] = e = {"fmap": fm[n], "ctx": None, "hw": tensor(j)[1:]}
            got = self._arena_slot(fm[n], Tn, keep) if (pooled and keys[j] is not None) else None
            if got is not None:
                e["arena"], e["slot"] = got
                e["fmap"] = got[0].put_fmap(got[1], fm[n])
            keep.append(e)
        for n, j in enumerate(uniq_c):
            e = new.get(j) or ent[j]
            e["ctx"] = e["arena"].put_ctx(e["slot"], ctx[0][n], ctx[1][n]) if "arena" in e else (ctx[0][n], ctx[1][n])
        self.cache_stats["fnet_runs"] += len(uniq_f)
        self.cache_stats["cnet_runs"] += len(uniq_c)
        if not prefetch:
            self.cache_stats["fnet_hits"] += Tn - len(need_f)
            self.cache_stats["cnet_hits"] += len(centre) - len(need_c)
        for i in range(Tn):
            if ent[i] is None:
                ent[i] = new[first[keys[i] if keys[i] is not None else ("#", i)]]
            if keys[i] is not None and self.feature_cache_frames > 0:
                self._feat[keys[i]] = ent[i]
                self._feat.move_to_end(keys[i])
        while len(self._feat) > self.feature_cache_frames:
            _, old = self._feat.popitem(last=False)
            if "arena" in old:
                old["arena"].free.append(old["slot"])
        if cur is not None:
            # cached tensors are produced and consumed on two streams (prefetch / caller): the caching allocator must not hand
            # a dropped entry's memory to the other stream while this stream's reads are still queued
            for e in ent:
                for t in (e["fmap"],) + (tuple(e["ctx"]) if e["ctx"] is not None else ()):
                    if t.is_cuda:
                        t.record_stream(cur)
        return ent

    def clear_feature_cache(self):
        self._feat.clear()
        self._arenas.clear()

    def _arena_slot(self, fm, Tn, keep):
        """(arena, slot) for a new cached frame whose feature map is `fm`; None when the cache is too small to pool (< 2 windows).
        A full arena takes the slot of its least recently used frame (never one of `keep`, the current window's entries)."""
        if self.feature_cache_frames < 2 * Tn:
            return None
        akey = (tuple(fm.shape), fm.dtype, fm.device)
        arena = self._arenas.get(akey)
        if arena is None:
            entry_bytes = 2 * fm.numel() * fm.element_size()          # fmap + (net, inp): 256 + 128 + 128 channels
            cap = max(2 * Tn, min(self.feature_cache_frames, self.feature_cache_bytes // max(entry_bytes, 1)))
            arena = self._arenas[akey] = _FeatureArena(cap)
        if not arena.free:
            for k in list(self._feat):
                e = self._feat[k]
                if e.get("arena") is arena and not any(e is x for x in keep):
                    del self._feat[k]
                    arena.free.append(e["slot"])
                    break
        if not arena.free:
            return None
        return arena, arena.free.pop()

    def is_model_loaded(self):
        return self.model is not None

    def get_model_info(self):
        if self.model is None:
            return {"status": "not_loaded"}
        return {
            "status": "loaded",
            "model_path": getattr(self.cfg, 'model', 'unknown') if self.cfg else 'unknown',
            "dataset": self.dataset,
            "architecture": self.architecture.upper(),
            "variant": self.variant,
            "config": {
                "decoder_depth": getattr(self.cfg, 'decoder_depth', 'default'),
                "corr_levels": getattr(self.cfg, 'corr_levels', 'default'),
                "corr_radius": getattr(self.cfg, 'corr_radius', 'default'),
            },
            "fast_mode": self.fast_mode,
            "device": str(self.device),
        }

    def get_device(self):
        return self.device

    def set_eval_mode(self):
        pass   # the engine is inference-only

    def get_memory_usage(self):
        dev = torch.device(self.device)
        if dev.type == 'cuda':
            return {"allocated": torch.cuda.memory_allocated(dev) / 1024**2,
                    "cached": torch.cuda.memory_reserved(dev) / 1024**2,
                    "max_allocated": torch.cuda.max_memory_allocated(dev) / 1024**2}
        return {"message": "Memory tracking only available for CUDA devices"}
