"""CPU restatement of the MemFlow (MemFlowNet_skflow) online inference path.  TEST INFRASTRUCTURE ONLY.

Follows /root/reference/MemFlow:
  inference/inference_core_skflow.py:7-54          InferenceCore (clear_memory, step)
  inference/memory_manager_skflow.py:38-124        MemoryManager.match_memory / add_memory / compress_features (einsum path)
  inference/kv_memory_store.py:32-91               KeyValueMemoryStore.add / sieve_by_range
  core/Networks/MemFlowNet/memory_util.py:41-65    do_softmax (top_k=None)
  core/Networks/MemFlowNet/MemFlow.py:66-128,163-184   encode_features / encode_context / upsample_flow
  core/Networks/MemFlowNet/sk2.py:86-109,182-215   motion encoder and update block (..._Mem_skflow)
  core/Networks/MemFlowNet/cnn.py:5-56,117-203     ResidualBlock / BasicEncoder (fnet: instance norm, cnet: batch norm, eval)
  configs/sintel_memflownet.py:17-53               val_decoder_depth 15, mem_every 1, max/min mid-term frames 2/1,
                                                   train_avg_length (368*768//64)*3/2 = 6624
and, for the application's call, processing/memflow_inference_isolated.py:52-86 (core_compute_flow below).

The PC block, correlation pyramid / lookup and convex upsampling are the functions of oracle/mof.py
(MemFlowNet/sk2.py:8-32, corr.py, MemFlow.py:173-184 are the same arithmetic as their VideoFlow counterparts).

Pinned by tests/golden/memflow_e2e.npz (reference MemFlowNet + InferenceCore executed in the build container with the
flash-attention branch disabled, tests/golden/make_golden_memflow.py); only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg may import this module.
"""
import math

import torch
import torch.nn.functional as F

from . import mof as omof

ITERS = 15                                      # configs/sintel_memflownet.py:24 (val_decoder_depth)
TRAIN_AVG_LENGTH = (368 * 768 // 64) * 3 / 2    # configs/sintel_memflownet.py:28
MAX_MID_TERM_FRAMES, MIN_MID_TERM_FRAMES = 2, 1  # configs/sintel_memflownet.py:35-36


# ---------------------------------------------------------------------------- parameters
def _encoder_shapes(p, norm):
    """BasicEncoder(output_dim=256): cnn.py:117-175.  norm='batch' adds the BatchNorm entries (the downsample norm is
    registered twice: as `norm3` and as `downsample.1`, cnn.py:23-24,44-45)."""
    s = {}

    def bn(name, c):
        if norm == "batch":
            for k in ("weight", "bias", "running_mean", "running_var"):
                s[f"{name}.{k}"] = (c,)

    def conv(name, co, ci, k):
        s[f"{name}.weight"] = (co, ci, k, k)
        s[f"{name}.bias"] = (co,)

    conv(p + ".conv1", 64, 3, 7)
    bn(p + ".norm1", 64)
    cin = 64
    for li, (dim, stride) in enumerate(((64, 1), (96, 2), (128, 2)), 1):
        for bi in range(2):
            q = f"{p}.layer{li}.{bi}"
            st = stride if bi == 0 else 1
            conv(q + ".conv1", dim, cin, 3)
            conv(q + ".conv2", dim, dim, 3)
            bn(q + ".norm1", dim)
            bn(q + ".norm2", dim)
            if st != 1:
                conv(q + ".downsample.0", dim, cin, 1)
                bn(q + ".norm3", dim)
                bn(q + ".downsample.1", dim)
            cin = dim
    conv(p + ".conv2", 256, 128, 1)
    return s


def memflow_param_shapes():
    """Floating-point entries of the reference MemFlowNet state_dict (cnet/fnet = basicencoder, gma = 'GMA-SK2')."""
    s = {}
    s.update(_encoder_shapes("cnet", "batch"))
    s.update(_encoder_shapes("fnet", "instance"))
    e = "update_block.encoder"
    ps = omof._pcblock_shapes
    s.update(ps(e + ".convc1", 324, 256, (1, 15)))                 # sk2.py:89-90
    s.update(ps(e + ".convc2", 256, 192, (1, 15)))
    s[e + ".convf1.weight"] = (128, 2, 1, 1)
    s[e + ".convf1.bias"] = (128,)
    s.update(ps(e + ".convf2", 128, 64, (1, 15)))
    s.update(ps(e + ".conv", 256, 126, (1, 15)))                    # sk2.py:95
    s.update(ps("update_block.gru", 512, 128, (1, 7)))
    s.update(ps("update_block.flow_head", 128, 2, (1, 15)))
    s["update_block.mask.0.weight"] = (256, 128, 3, 3)
    s["update_block.mask.0.bias"] = (256,)
    s["update_block.mask.2.weight"] = (576, 256, 1, 1)
    s["update_block.mask.2.bias"] = (576,)
    s["update_block.aggregator.to_v.weight"] = (128, 128, 1, 1)
    s["update_block.aggregator.gamma"] = (1,)
    s["att.to_qk.weight"] = (256, 128, 1, 1)
    s["att.pos_emb.rel_height.weight"] = (319, 128)
    s["att.pos_emb.rel_width.weight"] = (319, 128)
    return s


def memflow_synthetic_weights(seed=0, gain=1.18, flow_gain=0.33):
    """Seeded synthetic weights (checkpoints are absent), recipe of mof_synthetic_weights.  BatchNorm running
    statistics are made non-trivial (mean ~ U(+-0.1), var ~ 1 + U(0, 0.5)); mask head x100 as for BOF; the encoder
    convolutions get gain 2 so that eight ReLU layers keep a usable dynamic range; to_qk x8 so the read-out
    softmax is not flat.  With these defaults (update-block gain 1.18) the 15-iteration recurrence is well
    conditioned (a 1e-6 relative input perturbation moves the output by 2e-5 px, |flow| up to 10 px) while the
    stored memory frame changes the flow by up to 0.8 px, so the fixture really exercises the read-out.  (Gain 1.2 /
    flow gain 0.35 doubles both the flow range and the sensitivity: the fp16 tier then sits at the 5e-2 px bound.)"""
    shapes = memflow_param_shapes()
    enc = lambda k: 2.0 if (k.startswith("fnet.") or k.startswith("cnet.")) else gain
    ov = {
        "update_block.aggregator.gamma": lambda t: torch.full_like(t, 0.5),
        "update_block.flow_head.ffn2.2.weight": lambda t: t * flow_gain,
        "update_block.flow_head.ffn2.2.bias": lambda t: t * flow_gain,
        "update_block.mask.2.weight": lambda t: t * 100.0,
        "update_block.mask.2.bias": lambda t: t * 100.0,
        "att.to_qk.weight": lambda t: t * 8.0,
    }
    for k in shapes:
        if k.endswith("running_var"):
            ov[k] = lambda t: 1.0 + 2.0 * t.abs()
        elif k.endswith("running_mean"):
            ov[k] = lambda t: 0.8 * t
    sd = omof.synthetic_weights(shapes, seed, enc, ov)
    d = torch.arange(160).view(1, -1) - torch.arange(160).view(-1, 1)
    sd["att.pos_emb.rel_ind"] = d + 159
    for k in list(shapes):
        if ".downsample.1." in k:           # the same module as norm3 (cnn.py:44-45): a real checkpoint holds equal tensors
            sd[k] = sd[k.replace(".downsample.1.", ".norm3.")].clone()
    for k in list(shapes):
        if k.endswith("running_mean"):
            sd[k[:-len("running_mean")] + "num_batches_tracked"] = torch.tensor(0, dtype=torch.long)
    return sd


# ---------------------------------------------------------------------------- encoders (boundary: library convolutions)
def _norm(x, sd, name, norm):
    if norm == "batch":                                     # eval-mode BatchNorm2d
        return F.batch_norm(x, sd[name + ".running_mean"], sd[name + ".running_var"], sd[name + ".weight"], sd[name + ".bias"],
                            False, 0.0, 1e-5)
    return F.instance_norm(x, eps=1e-5)                     # InstanceNorm2d default: no affine, no running stats


def _res_block(x, sd, q, norm, stride):
    """ResidualBlock.forward, cnn.py:48-56"""
    y = F.relu(_norm(F.conv2d(x, sd[q + ".conv1.weight"], sd[q + ".conv1.bias"], stride=stride, padding=1), sd, q + ".norm1", norm))
    y = F.relu(_norm(F.conv2d(y, sd[q + ".conv2.weight"], sd[q + ".conv2.bias"], padding=1), sd, q + ".norm2", norm))
    if stride != 1:
        x = _norm(F.conv2d(x, sd[q + ".downsample.0.weight"], sd[q + ".downsample.0.bias"], stride=stride), sd, q + ".norm3", norm)
    return F.relu(x + y)


def basic_encoder(x, sd, p, norm):
    """BasicEncoder.forward (eval), cnn.py:177-203: [N,3,H,W] -> [N,256,H/8,W/8]"""
    x = F.relu(_norm(F.conv2d(x, sd[p + ".conv1.weight"], sd[p + ".conv1.bias"], stride=2, padding=3), sd, p + ".norm1", norm))
    for li, stride in ((1, 1), (2, 2), (3, 2)):
        x = _res_block(x, sd, f"{p}.layer{li}.0", norm, stride)
        x = _res_block(x, sd, f"{p}.layer{li}.1", norm, 1)
    return F.conv2d(x, sd[p + ".conv2.weight"], sd[p + ".conv2.bias"])


def encode_context(frame, sd):
    """MemFlowNet.encode_context, MemFlow.py:100-128 (4-D input): -> query, key, net, inp [B,128,h,w]"""
    cnet = basic_encoder(frame, sd, "cnet", "batch")
    net, inp = torch.split(cnet, [128, 128], dim=1)
    net, inp = torch.tanh(net), torch.relu(inp)
    q, k = F.conv2d(inp, sd["att.to_qk.weight"]).chunk(2, dim=1)
    return q, k, net, inp


def encode_features(frames, sd):
    """MemFlowNet.encode_features, MemFlow.py:66-98 (5-D input [B,T,3,H,W]) -> fmaps [B,T,256,h,w] fp32"""
    b, t = frames.shape[:2]
    f = basic_encoder(frames.flatten(0, 1), sd, "fnet", "instance").float()
    return f.view(b, t, *f.shape[-3:])


# ---------------------------------------------------------------------------- update block
def motion_encoder(flow, corr, sd, p="update_block.encoder"):
    """SKMotionEncoder6_Deep_nopool_res_Mem_skflow.forward, sk2.py:97-109"""
    k = (1, 15)
    cor = omof.pcblock(F.gelu(omof.pcblock(corr, sd, p + ".convc1", k)), sd, p + ".convc2", k)
    flo = omof.pcblock(omof._conv(flow, sd, p + ".convf1"), sd, p + ".convf2", k)
    out = omof.pcblock(torch.cat([cor, flo], 1), sd, p + ".conv", k)
    return torch.cat([out, flow], 1)


def update_block(net, inp, mf, mfg, sd, with_mask=True):
    """SKUpdateBlock6_Deep_nopoolres_AllDecoder2_Mem_skflow.forward, sk2.py:205-215"""
    p = "update_block"
    net = omof.pcblock(torch.cat([net, inp, mf, mfg], 1), sd, p + ".gru", (1, 7))
    delta = omof.pcblock(net, sd, p + ".flow_head", (1, 15))
    mask = None
    if with_mask:
        mask = 0.25 * omof._conv(F.relu(omof._conv(net, sd, p + ".mask.0", padding=1)), sd, p + ".mask.2")
    return net, mask, delta


# ---------------------------------------------------------------------------- memory
def memory_readout(query, keys, values, train_avg_length=TRAIN_AVG_LENGTH):
    """match_memory on flattened tensors (memory_manager_skflow.py:38-97, einsum branch): query [B,128,P],
    keys [B,128,T], values [B,128,T] -> [B,128,P].  scale = 128^-1/2 * log(T) / log(train_avg_length) (:57);
    softmax over the T memory elements (do_softmax, memory_util.py:57-61)."""
    scale = query.shape[1] ** -0.5 * math.log(keys.shape[-1], train_avg_length)
    sim = torch.einsum("bcl,bct->btl", query, keys) * scale
    aff = torch.softmax(sim, dim=1)
    return values @ aff


class InferenceCore:
    """inference_core_skflow.py:7-54 with the MemoryManager / KeyValueMemoryStore state inlined: the working memory
    is a pair of [1,128,n*P] tensors; when it reaches max_mid_term_frames*P elements it is cut back to the last
    min_mid_term_frames*P (memory_manager_skflow.py:117-124, kv_memory_store.py:62-76)."""

    def __init__(self, sd, iters=ITERS, mem_every=1, train_avg_length=TRAIN_AVG_LENGTH, max_frames=MAX_MID_TERM_FRAMES,
                 min_frames=MIN_MID_TERM_FRAMES):
        self.sd, self.iters, self.mem_every, self.tal = sd, iters, mem_every, train_avg_length
        self.max_frames, self.min_frames = max_frames, min_frames
        self.clear_memory()

    def clear_memory(self):
        self.curr_ti = -1
        self.last_mem_ti = -self.mem_every
        self.mem_k = self.mem_v = None

    def step(self, images, end=False, flow_init=None):
        """images [1,2,3,H,W] in [-1,1] (H, W multiples of 8) -> (flow_low [1,2,h,w], flow_up [1,2,H,W])"""
        sd = self.sd
        self.curr_ti += 1
        is_mem_frame = (self.curr_ti - self.last_mem_ti >= self.mem_every) and not end
        query, key, net, inp = encode_context(images[:, 0], sd)
        fmaps = encode_features(images, sd)
        B, _, h, w = query.shape
        coords0 = omof.pixel_grid(B, h, w, query.device)
        coords1 = coords0.clone() if flow_init is None else coords0 + flow_init
        pyr = omof.corr_pyramid(fmaps[:, 0], fmaps[:, 1])
        qf, kf = query.flatten(2), key.flatten(2)
        keys = kf if self.mem_k is None else torch.cat([self.mem_k, kf], -1)
        mask = value = None
        for it in range(self.iters):
            corr = omof.corr_lookup(pyr, coords1)
            mf = motion_encoder(coords1 - coords0, corr, sd)
            value = omof._conv(mf, sd, "update_block.aggregator.to_v")
            vf = value.flatten(2)
            values = vf if self.mem_v is None else torch.cat([self.mem_v, vf], -1)
            readout = memory_readout(qf, keys, values, self.tal).view(B, -1, h, w)
            mfg = mf + sd["update_block.aggregator.gamma"] * readout
            net, mask, delta = update_block(net, inp, mf, mfg, sd, with_mask=(it == self.iters - 1))
            coords1 = coords1 + delta
        flow_up = omof.convex_upsample8(coords1 - coords0, mask)
        if is_mem_frame:
            vf = value.flatten(2)
            self.mem_k = kf if self.mem_k is None else torch.cat([self.mem_k, kf], -1)
            self.mem_v = vf if self.mem_v is None else torch.cat([self.mem_v, vf], -1)
            P = h * w
            if self.mem_k.shape[-1] >= self.max_frames * P:
                self.mem_k = self.mem_k[:, :, -self.min_frames * P:]
                self.mem_v = self.mem_v[:, :, -self.min_frames * P:]
            self.last_mem_ti = self.curr_ti
        return coords1 - coords0, flow_up


def normalize_frames(x):
    """The range heuristic of processing/memflow_inference_isolated.py:64-69."""
    m = float(x.max())
    if m > 2.0:
        return 2 * (x / 255.0) - 1.0
    if m > 1.0:
        return 2 * x - 1.0
    return x


def core_compute_flow(window, sd, iters=ITERS):
    """What MemFlowCore.compute_flow_from_tensor returns (processing/memflow_core.py:163-184 ->
    memflow_inference_isolated.py:52-86): normalise, replicate-pad to x8, a FRESH InferenceCore, one step on the
    last two frames with end=True (so nothing is ever memorised and the read-out sees the current frame only),
    unpad -> [2,H,W]."""
    H, W = window.shape[-2:]
    x = omof.pad_frames(normalize_frames(window.float()))
    _, up = InferenceCore(sd, iters).step(x[:, -2:], end=True)
    return omof.unpad(up[0], H, W)
