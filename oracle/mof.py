"""CPU restatement of the VideoFlow MOF (multi-frame) inference path (oracle; test infrastructure only).

Functional torch-fp32 restatement of the reference modules, operating on a flat
``state_dict`` with the upstream checkpoint key names.  Each function cites the
reference lines it follows (paths relative to the reference checkout,
``VideoFlow/core/Networks/MOFNetStack`` abbreviated ``MOF/``).  Pinned by
``tests/golden/mof_*.npz`` which hold outputs of the unmodified reference
modules (see ``tests/golden/make_golden.py``).
"""
import math

import torch
import torch.nn.functional as F

from .twins import twins_svt_forward, twins_param_shapes

CORR_LEVELS = 4
CORR_RADIUS = 4          # forced in MOF/network.py:23
ITERS = 32               # VideoFlow/configs/multiframes_sintel_submission.py:36


# ----------------------------------------------------------------------------
# padding, grids, sampling            VideoFlow/core/utils/utils.py
# ----------------------------------------------------------------------------
def pad_amounts(ht, wd, mode="sintel"):
    """(left, right, top, bottom) replicate padding to a multiple of 8. utils.py:9-18"""
    ph = (((ht // 8) + 1) * 8 - ht) % 8
    pw = (((wd // 8) + 1) * 8 - wd) % 8
    if mode == "sintel":
        return pw // 2, pw - pw // 2, ph // 2, ph - ph // 2
    return pw // 2, pw - pw // 2, 0, ph


def pad_frames(x, mode="sintel"):
    """x [..., H, W] -> replicate padded. utils.py:20-24 (called with a 5-D tensor at
    processing/videoflow_core.py:182-183, hence the trailing zero pair)."""
    l, r, t, b = pad_amounts(x.shape[-2], x.shape[-1], mode)
    if x.dim() == 5:
        return F.pad(x, [l, r, t, b, 0, 0], mode="replicate")
    return F.pad(x, [l, r, t, b], mode="replicate")


def unpad(x, ht, wd, mode="sintel"):
    """utils.py:26-29"""
    l, r, t, b = pad_amounts(ht, wd, mode)
    H, W = x.shape[-2:]
    return x[..., t:H - b, l:W - r]


def pixel_grid(n, h, w, device=None):
    """[n,2,h,w], channel 0 = x, channel 1 = y. utils.py:97-100"""
    ys, xs = torch.meshgrid(torch.arange(h, device=device), torch.arange(w, device=device), indexing="ij")
    return torch.stack([xs, ys], 0).float()[None].repeat(n, 1, 1, 1)


def sample_px(img, coords):
    """Bilinear sample ``img`` [N,C,H,W] at pixel coords [N,h,w,2] (x,y); zeros outside,
    align_corners=True. utils.py:62-76"""
    H, W = img.shape[-2:]
    gx = 2 * coords[..., 0:1] / (W - 1) - 1
    gy = 2 * coords[..., 1:2] / (H - 1) - 1
    return F.grid_sample(img, torch.cat([gx, gy], -1), align_corners=True)


# ----------------------------------------------------------------------------
# correlation                          MOF/corr.py:59-107
# ----------------------------------------------------------------------------
def corr_volume(f1, f2):
    """All-pairs correlation. f1,f2 [B,C,h,w] -> [B,h,w,1,h,w]. corr.py:99-107"""
    B, C, h, w = f1.shape
    v = torch.matmul(f1.reshape(B, C, h * w).transpose(1, 2), f2.reshape(B, C, h * w))
    return v.view(B, h, w, 1, h, w) / math.sqrt(C)


def corr_pyramid(f1, f2, levels=CORR_LEVELS):
    """List of [B*h*w,1,h_l,w_l]; each level is avg_pool2d(2,2) of the previous. corr.py:60-74"""
    v = corr_volume(f1, f2)
    B, h, w = v.shape[:3]
    v = v.reshape(B * h * w, 1, h, w)
    pyr = [v]
    for _ in range(levels - 1):
        v = F.avg_pool2d(v, 2, stride=2)
        pyr.append(v)
    return pyr


def corr_lookup(pyr, coords, radius=CORR_RADIUS):
    """coords [B,2,h,w] -> [B, L*(2r+1)^2, h, w]; channel l*81 + i*9 + j samples level l at
    (x/2^l + i - r, y/2^l + j - r) (x offset is the slow index). corr.py:76-97"""
    r = radius
    n = 2 * r + 1
    c = coords.permute(0, 2, 3, 1)
    B, h, w, _ = c.shape
    d = torch.linspace(-r, r, n, device=coords.device)
    # stack(meshgrid(dy, dx)) added to (x, y): first grid (varies along dim 0) goes to x
    off = torch.stack(torch.meshgrid(d, d, indexing="ij"), -1).view(1, n, n, 2)
    out = []
    for l, vol in enumerate(pyr):
        centre = c.reshape(B * h * w, 1, 1, 2) / 2 ** l
        s = sample_px(vol, centre + off)
        out.append(s.view(B, h, w, -1))
    return torch.cat(out, -1).permute(0, 3, 1, 2).contiguous().float()


def alt_corr_forward(fmap1, fmap2, coords, r):
    """alt_cuda_corr.forward restated (VideoFlow/alt_cuda_corr/correlation_kernel.cu:18-119): fmap1 [B,H1,W1,C],
    fmap2 [B,H2,W2,C] NHWC, coords [B,N,H1,W1,2] (x, y) -> corr [B,N,(2r+1)^2,H1,W1].  For every query the kernel
    forms s[iy][ix] = <f1, f2[floor(y)-r+iy, floor(x)-r+ix]> (0 out of bounds, :66-87) on the (2r+2)^2 integer grid
    and scatters it with the weights nw = dy*dx, ne = dy*(1-dx), sw = (1-dy)*dx, se = (1-dy)*(1-dx) into channel
    iy' + (2r+1)*ix' (:89-112), i.e. an x-major tap order and no 1/sqrt(C) factor.
    PARITY UNPINNED directly (the extension needs a GPU build to run); anchored on its documented equivalence with
    CorrBlock (VideoFlow/README.md:112, corr.py:149-177 vs :59-97), which tests/test_oracle_golden.py checks."""
    B, H1, W1, C = fmap1.shape
    H2, W2 = fmap2.shape[1:3]
    N = coords.shape[1]
    rd = 2 * r + 1
    x, y = coords[..., 0].double(), coords[..., 1].double()              # [B,N,H1,W1]
    fx, fy = torch.floor(x), torch.floor(y)
    dx, dy = (x - fx), (y - fy)
    f1 = fmap1.double()
    f2 = fmap2.double().reshape(B, H2 * W2, C)
    s = torch.zeros(B, N, rd + 1, rd + 1, H1, W1, dtype=torch.float64)
    for iy in range(rd + 1):
        for ix in range(rd + 1):
            h2 = (fy - r + iy).long()
            w2 = (fx - r + ix).long()
            ok = (h2 >= 0) & (h2 < H2) & (w2 >= 0) & (w2 < W2)
            idx = (h2.clamp(0, H2 - 1) * W2 + w2.clamp(0, W2 - 1)).reshape(B, N * H1 * W1)
            g = torch.gather(f2, 1, idx.unsqueeze(-1).expand(-1, -1, C)).reshape(B, N, H1, W1, C)
            s[:, :, iy, ix] = (g * f1.unsqueeze(1)).sum(-1) * ok
    out = torch.zeros(B, N, rd * rd, H1, W1, dtype=torch.float64)
    for ix in range(rd):
        for iy in range(rd):
            out[:, :, iy + rd * ix] = ((1 - dy) * (1 - dx) * s[:, :, iy, ix] + (1 - dy) * dx * s[:, :, iy, ix + 1]
                                       + dy * (1 - dx) * s[:, :, iy + 1, ix] + dy * dx * s[:, :, iy + 1, ix + 1])
    return out.float()


def alt_corr_block(fmap1, fmap2, coords, num_levels=CORR_LEVELS, radius=CORR_RADIUS):
    """AlternateCorrBlock(fmap1, fmap2)(coords), MOFNetStack/corr.py:149-177: fmap [B,C,h,w], coords [B,2,h,w]."""
    c = coords.permute(0, 2, 3, 1)
    B, H, W, _ = c.shape
    f1 = fmap1.permute(0, 2, 3, 1).contiguous()
    out = []
    for i in range(num_levels):
        f2 = fmap2.permute(0, 2, 3, 1).contiguous()
        out.append(alt_corr_forward(f1, f2, (c / 2 ** i).reshape(B, 1, H, W, 2), radius).squeeze(1))
        fmap2 = F.avg_pool2d(fmap2, 2, stride=2)
    return torch.stack(out, 1).reshape(B, -1, H, W) / torch.sqrt(torch.tensor(fmap1.shape[1]).float())


# ----------------------------------------------------------------------------
# update block                         MOF/stack.py, MOF/gma.py
# ----------------------------------------------------------------------------
def _conv(x, sd, p, **kw):
    return F.conv2d(x, sd[p + ".weight"], sd.get(p + ".bias"), **kw)


def pcblock(x, sd, p, k_conv):
    """PCBlock4_Deep_nopool_res. stack.py:8-32"""
    C = x.shape[1]
    x = F.gelu(x + _conv(F.gelu(_conv(x, sd, p + ".ffn1.0")), sd, p + ".ffn1.2"))
    for i, k in enumerate(k_conv):
        x = F.gelu(x + _conv(x, sd, f"{p}.conv_list.{i}", padding=k // 2, groups=C))
    x = F.gelu(x + _conv(x, sd, p + ".pw"))
    return _conv(F.gelu(_conv(x, sd, p + ".ffn2.0")), sd, p + ".ffn2.2")


def gma_attention(inp, sd, p="att"):
    """Attention.forward, 1 head of 128. gma.py:54-76 -> [B,1,P,P]"""
    B, C, h, w = inp.shape
    qk = _conv(inp, sd, p + ".to_qk")
    q, k = qk.chunk(2, dim=1)
    q = q.reshape(B, C, h * w).transpose(1, 2) * (C ** -0.5)
    k = k.reshape(B, C, h * w)
    return torch.softmax(q @ k, dim=-1)[:, None]


def gma_aggregate(attn, fmap, sd, p="update_block.aggregator"):
    """Aggregate.forward (dim == inner_dim so no project). gma.py:102-115"""
    B, C, h, w = fmap.shape
    v = _conv(fmap, sd, p + ".to_v").reshape(B, C, h * w).transpose(1, 2)
    o = (attn[:, 0] @ v).transpose(1, 2).reshape(B, C, h, w)
    return fmap + sd[p + ".gamma"] * o


def motion_encoder(mhs, ffl, bfl, coords0, fcorr, bcorr, bs, sd, p="update_block.encoder"):
    """SKMotionEncoder6_Deep_nopool_res.forward. stack.py:69-103"""
    BN, _, H, W = ffl.shape
    N = BN // bs
    if mhs is None:
        mhs = sd[p + ".init_hidden_state"].repeat(bs, N, 1, H, W)
    else:
        mhs = mhs.reshape(bs, N, -1, H, W)
    z = torch.zeros(bs, 1, 48, H, W, device=ffl.device)
    nxt = torch.cat([mhs[:, 1:], z], 1).reshape(BN, -1, H, W)       # state of frame t+1
    prv = torch.cat([z, mhs[:, :N - 1]], 1).reshape(BN, -1, H, W)   # state of frame t-1
    fw = sample_px(nxt.float(), (ffl + coords0).permute(0, 2, 3, 1))
    bw = sample_px(prv.float(), (bfl + coords0).permute(0, 2, 3, 1))
    k = (1, 15)
    fc = pcblock(fcorr, sd, p + ".convc1", k)
    bc = pcblock(bcorr, sd, p + ".convc1", k)
    cor = pcblock(F.gelu(torch.cat([fc, bc], 1)), sd, p + ".convc2", k)
    flow = torch.cat([ffl, bfl], 1)
    flo = pcblock(_conv(flow, sd, p + ".convf1_"), sd, p + ".convf2", k)
    x = torch.cat([cor, flo, fw, bw, mhs.reshape(BN, -1, H, W)], 1)
    out = pcblock(x, sd, p + ".conv", k)
    return torch.cat([out[:, :124], flow], 1), out[:, 124:172]


def update_block(net, mhs, inp, fcorr, bcorr, ffl, bfl, coords0, attn, bs, sd, with_mask=True):
    """SKUpdateBlock6_Deep_nopoolres_AllDecoder2.forward. stack.py:127-140"""
    p = "update_block"
    mf, mhs = motion_encoder(mhs, ffl, bfl, coords0, fcorr, bcorr, bs, sd)
    mfg = gma_aggregate(attn, mf, sd)
    net = pcblock(torch.cat([net, inp, mf, mfg], 1), sd, p + ".gru", (1, 7))
    delta = pcblock(net, sd, p + ".flow_head", (1, 15))
    mask = None
    if with_mask:
        m = F.relu(_conv(net, sd, p + ".mask.0", padding=1))
        mask = 100.0 * _conv(m, sd, p + ".mask.2")
    return net, mhs, mask, delta


def convex_upsample8(flow, mask):
    """flow [N,2,h,w], mask [N,576,h,w] -> [N,2,8h,8w]. MOF/network.py:97-108"""
    N, _, H, W = flow.shape
    m = torch.softmax(mask.view(N, 1, 9, 8, 8, H, W), dim=2)
    u = F.unfold(8 * flow, [3, 3], padding=1).view(N, 2, 9, 1, 1, H, W)
    u = torch.sum(m * u, dim=2).permute(0, 1, 4, 2, 5, 3)
    return u.reshape(N, 2, 8 * H, 8 * W)


# ----------------------------------------------------------------------------
# network                              MOF/network.py:138-215
# ----------------------------------------------------------------------------
def mof_encode(images, sd, encoder=None, amp=None):
    """images [1,T,3,H,W] (raw caller values) -> fmaps [T,256,h,w], net/inp [T-2,128,h,w].
    network.py:144-170.  ``encoder(x, prefix)`` overrides the Twins restatement."""
    B, T, _, H, W = images.shape
    x = 2 * (images / 255.0) - 1.0
    enc = encoder or (lambda im, pre: twins_svt_forward(im, sd, pre + ".svt"))
    with _amp(amp, images.device.type):          # network.py:149,160: both encoders run under autocast on CUDA
        fmaps = enc(x.reshape(B * T, 3, H, W), "fnet").float()
        cnet = enc(x[:, 1:T - 1].reshape(B * (T - 2), 3, H, W), "cnet")
        net, inp = torch.split(cnet, [128, 128], dim=1)
        return fmaps, torch.tanh(net), torch.relu(inp)


def _amp(dtype, device_type="cpu"):
    """Emulation of the reference's CUDA mixed-precision path (network.py:13,149,160,191: encoders, attention and
    the update block under fp16 autocast; volume build and lookup in fp32, network.py:151 / corr.py:97) with
    torch's CPU autocast.  Used only to measure the reference's OWN reduced-precision deviation (the noise floor
    the fp16 tier is compared with, scripts/fp16_floor.py); dtype=None is the fp32 path."""
    import contextlib
    return torch.autocast(device_type, dtype=dtype) if dtype is not None else contextlib.nullcontext()


def mof_refine(fmaps, net, inp, sd, iters=ITERS, all_iters=False, trace=None, amp=None):
    """Correlation build + ``iters`` update iterations + convex upsampling.
    fmaps [T,256,h,w] -> flows [1, 2(T-2), 2, 8h, 8w] (forward flows then backward flows).
    network.py:151-215 with B=1."""
    T = fmaps.shape[0]
    Fc = T - 2
    h, w = fmaps.shape[-2:]
    fpyr = corr_pyramid(fmaps[1:T - 1], fmaps[2:T])
    bpyr = corr_pyramid(fmaps[1:T - 1], fmaps[0:T - 2])
    dt_ = fmaps.device.type
    with _amp(amp, dt_):
        attn = gma_attention(inp, sd)
    c0 = pixel_grid(Fc, h, w, fmaps.device)
    fc1, bc1 = c0.clone(), c0.clone()
    mhs = None
    outs = []
    for it in range(iters):
        fcorr = corr_lookup(fpyr, fc1)
        bcorr = corr_lookup(bpyr, bc1)
        last = it == iters - 1
        with _amp(amp, dt_):
            net, mhs, mask, delta = update_block(net, mhs, inp, fcorr, bcorr, fc1 - c0, bc1 - c0, c0, attn, 1, sd,
                                                 with_mask=(last or all_iters))
        fc1 = fc1 + delta[:, 0:2]
        bc1 = bc1 + delta[:, 2:4]
        if trace is not None:
            trace.append((fc1 - c0, bc1 - c0))
        if last or all_iters:
            fu = convex_upsample8((fc1 - c0).float(), mask[:, :576].float())
            bu = convex_upsample8((bc1 - c0).float(), mask[:, 576:].float())
            outs.append(torch.cat([fu[None], bu[None]], 1))
    return outs if all_iters else outs[-1]


def mof_forward(images, sd, iters=ITERS, encoder=None):
    fmaps, net, inp = mof_encode(images, sd, encoder)
    return mof_refine(fmaps, net, inp, sd, iters)


def core_compute_flow(window, sd, iters=ITERS, encoder=None):
    """VideoFlowCore.compute_flow_from_tensor: pad, forward, unpad, take index T-2
    (first backward flow). processing/videoflow_core.py:181-198"""
    H, W = window.shape[-2:]
    flows = mof_forward(pad_frames(window), sd, iters, encoder)
    flows = unpad(flows, H, W)
    return flows[0, flows.shape[1] // 2]


# ----------------------------------------------------------------------------
# host logic                           processing/videoflow_processor.py
# ----------------------------------------------------------------------------
def window_indices(n_frames, idx, T):
    """Frame indices of the centred window, padded by repeating the first frame when the
    window is clipped at the clip start, otherwise the last. videoflow_processor.py:134-150"""
    half = T // 2
    s = max(0, idx - half)
    e = min(n_frames, idx + half + 1)
    ids = list(range(s, e))
    while len(ids) < T:
        if s == 0:
            ids.insert(0, ids[0])
        else:
            ids.append(ids[-1])
    return ids[:T]


def frames_to_window(frames, idx, T):
    """uint8 HWC frames -> [1,T,3,H,W] fp32 in [0,1]. videoflow_processor.py:152-162"""
    import numpy as np
    ts = []
    for i in window_indices(len(frames), idx, T):
        f = frames[i]
        a = f.astype(np.float32) / 255.0 if f.dtype == np.uint8 else f.astype(np.float32)
        ts.append(torch.from_numpy(a).permute(2, 0, 1))
    return torch.stack(ts)[None]


def tile_grid(width, height, tile=1280):
    """Disjoint fixed-size tiles, ragged at the right/bottom edge. videoflow_processor.py:73-110"""
    cols = -(-width // tile)
    rows = -(-height // tile)
    info = []
    for r in range(rows):
        for c in range(cols):
            x, y = c * tile, r * tile
            info.append({"x": x, "y": y, "width": min(tile, width - x), "height": min(tile, height - y),
                         "col": c, "row": r})
    return tile, tile, cols, rows, info


# ----------------------------------------------------------------------------
# parameter manifest + seeded synthetic weights
# ----------------------------------------------------------------------------
def _pcblock_shapes(p, cin, cout, k_conv):
    hid = int(1.5 * cin)
    out = {}
    for i, k in enumerate(k_conv):
        out[f"{p}.conv_list.{i}.weight"] = (cin, 1, k, k)
        out[f"{p}.conv_list.{i}.bias"] = (cin,)
    out[f"{p}.ffn1.0.weight"] = (hid, cin, 1, 1)
    out[f"{p}.ffn1.0.bias"] = (hid,)
    out[f"{p}.ffn1.2.weight"] = (cin, hid, 1, 1)
    out[f"{p}.ffn1.2.bias"] = (cin,)
    out[f"{p}.pw.weight"] = (cin, cin, 1, 1)
    out[f"{p}.pw.bias"] = (cin,)
    out[f"{p}.ffn2.0.weight"] = (hid, cin, 1, 1)
    out[f"{p}.ffn2.0.bias"] = (hid,)
    out[f"{p}.ffn2.2.weight"] = (cout, hid, 1, 1)
    out[f"{p}.ffn2.2.bias"] = (cout,)
    return out


def mof_param_shapes():
    """Every floating-point entry of the reference MOFNet ``state_dict`` (checked against the
    real module in tests/golden/make_golden.py).  Includes the dead weights that strict
    checkpoint loading needs.  The int64 buffer ``att.pos_emb.rel_ind`` is
    added by :func:`mof_synthetic_weights`."""
    s = {}
    s.update(twins_param_shapes("fnet.svt"))
    s.update(twins_param_shapes("cnet.svt"))
    e = "update_block.encoder"
    s.update(_pcblock_shapes(e + ".convc1", 324, 128, (1, 15)))
    s.update(_pcblock_shapes(e + ".convc2", 256, 192, (1, 15)))
    s[e + ".init_hidden_state"] = (1, 1, 48, 1, 1)
    s[e + ".convf1_.weight"] = (128, 4, 1, 1)
    s[e + ".convf1_.bias"] = (128,)
    s.update(_pcblock_shapes(e + ".convf2", 128, 64, (1, 15)))
    s.update(_pcblock_shapes(e + ".conv", 400, 172, (1, 15)))
    for i, (ci, co) in zip((0, 2, 4), ((214, 64), (64, 64), (64, 43))):
        s[f"{e}.velocity_update_block.mlp.{i}.weight"] = (co, ci, 3, 3)
        s[f"{e}.velocity_update_block.mlp.{i}.bias"] = (co,)
    s.update(_pcblock_shapes("update_block.gru", 512, 128, (1, 7)))
    s.update(_pcblock_shapes("update_block.flow_head", 128, 4, (1, 15)))
    s["update_block.mask.0.weight"] = (256, 128, 3, 3)
    s["update_block.mask.0.bias"] = (256,)
    s["update_block.mask.2.weight"] = (1152, 256, 1, 1)
    s["update_block.mask.2.bias"] = (1152,)
    s["update_block.aggregator.to_v.weight"] = (128, 128, 1, 1)
    s["update_block.aggregator.gamma"] = (1,)
    s["att.to_qk.weight"] = (256, 128, 1, 1)
    s["att.pos_emb.rel_height.weight"] = (319, 128)
    s["att.pos_emb.rel_width.weight"] = (319, 128)
    return s


def synthetic_weights(shapes, seed=0, gain=None, overrides=None):
    """Deterministic (numpy PCG64) weights: U(+-g/sqrt(fan_in)) for weights and biases with
    ``g = gain(key)`` (default 1), LayerNorm weights = 1 + noise.  Platform independent, so
    the GPU box regenerates exactly what the golden fixtures were produced with."""
    import numpy as np
    rng = np.random.default_rng(seed)
    sd = {}
    for k in sorted(shapes):
        shp = shapes[k]
        fan_in = int(np.prod(shp[1:])) if len(shp) >= 2 else 64
        b = (gain(k) if gain else 1.0) / math.sqrt(fan_in)
        a = rng.uniform(-b, b, size=shp).astype(np.float32)
        if ".norm" in k and k.endswith(".weight") and len(shp) == 1:
            a = (1.0 + a).astype(np.float32)
        sd[k] = torch.from_numpy(a)
    for k, f in (overrides or {}).items():
        sd[k] = f(sd[k])
    return sd


DRIFT = (0.6, 0.25, -0.5, -0.3)     # px per iteration at 1/8 resolution: fwd (x, y), bwd (x, y)


def mof_synthetic_weights(seed=0, gain=1.4, flow_gain=0.25, drift=None):
    """Seeded synthetic MOFNet weights (checkpoints are absent: .MISSING_LARGE_BLOBS).
    ``drift`` (4 floats, e.g. ``DRIFT``) is added to ``update_block.flow_head.ffn2.2.bias``: every iteration then
    moves the forward / backward coordinates by about that many low-resolution texels, so over 32 iterations
    the lookups travel ~19 texels (150 px), through the border and out of range, while the recurrence keeps the
    conditioning of the gain it is combined with (the "large-motion" fixtures, tests/golden/make_golden_drift.py).
    Update-block weights get ``gain`` 1.4 so the recurrence is input-sensitive (low-res flow
    std ~2 px, |flow| up to ~12 px after 32 iterations on the float test clip: in-range, border and out-of-range
    taps all occur) without diverging; ``aggregator.gamma`` is 0.5 (it initialises to 0,
    gma.py:95, which would hide the aggregate path); the last flow-head layer is damped."""
    enc = lambda k: 1.0 if (k.startswith("fnet.") or k.startswith("cnet.")) else gain
    sd = synthetic_weights(mof_param_shapes(), seed, enc, {
        "update_block.aggregator.gamma": lambda t: torch.full_like(t, 0.5),
        "update_block.flow_head.ffn2.2.weight": lambda t: t * flow_gain,
        "update_block.flow_head.ffn2.2.bias": lambda t: t * flow_gain,
    })
    if drift is not None:
        sd["update_block.flow_head.ffn2.2.bias"] = sd["update_block.flow_head.ffn2.2.bias"] + torch.tensor(drift, dtype=torch.float32)
    d = torch.arange(160).view(1, -1) - torch.arange(160).view(-1, 1)
    sd["att.pos_emb.rel_ind"] = d + 159            # gma.py:16-18 (buffer, unused in forward)
    return sd
