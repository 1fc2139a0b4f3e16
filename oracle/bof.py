"""CPU restatement of the VideoFlow BOF (3-frame, bidirectional) inference path.  TEST INFRASTRUCTURE ONLY.

Follows /root/reference/VideoFlow/core/Networks/BOFNet/network.py:82-145 (BOFNet.forward) and
BOFNet/sk2.py:33-92 (SKMotionEncoder6_Deep_nopool_res without hidden state,
SKUpdateBlock6_Deep_nopoolres_AllDecoder2 with the 0.25 mask scale).  The shared pieces (PC block,
correlation pyramid / lookup, GMA attention and aggregate, convex upsampling, Twins stages 0-1) are
the functions of oracle/mof.py: BOFNet/corr.py:12-61, BOFNet/gma.py and BOFNet/network.py:68-80 are
line-for-line the same arithmetic as their MOFNetStack counterparts.

Pinned by tests/golden/bof_e2e.npz (reference BOFNet executed in the build container by
tests/golden/make_golden_bof.py); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
leg may import this module.
"""
import torch
import torch.nn.functional as F

from . import mof as omof

ITERS = 12      # VideoFlow/configs/things.py:31 (BOFNet.decoder_depth)


def bof_param_shapes():
    """Floating-point entries of the reference BOFNet state_dict (cnet/fnet = twins, gma = 'GMA-SK2')."""
    s = {}
    s.update(omof.twins_param_shapes("fnet.svt"))
    s.update(omof.twins_param_shapes("cnet.svt"))
    e = "update_block.encoder"
    ps = omof._pcblock_shapes
    s.update(ps(e + ".convc1", 324, 128, (1, 15)))
    s.update(ps(e + ".convc2", 256, 192, (1, 15)))
    s[e + ".convf1_.weight"] = (128, 4, 1, 1)
    s[e + ".convf1_.bias"] = (128,)
    s.update(ps(e + ".convf2", 128, 64, (1, 15)))
    s.update(ps(e + ".conv", 256, 124, (1, 15)))                   # sk2.py:43
    s.update(ps("update_block.gru", 512, 128, (1, 7)))
    s.update(ps("update_block.flow_head", 128, 4, (1, 15)))
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


def bof_synthetic_weights(seed=0, gain=1.0, flow_gain=0.25):
    """Seeded synthetic BOFNet weights (checkpoints are absent), same recipe as mof_synthetic_weights.
    The mask head is scaled x100 so that 0.25 * mask (sk2.py:91) spans the same softmax range as MOF's."""
    enc = lambda k: 1.0 if (k.startswith("fnet.") or k.startswith("cnet.")) else gain
    sd = omof.synthetic_weights(bof_param_shapes(), seed, enc, {
        "update_block.aggregator.gamma": lambda t: torch.full_like(t, 0.5),
        "update_block.flow_head.ffn2.2.weight": lambda t: t * flow_gain,
        "update_block.flow_head.ffn2.2.bias": lambda t: t * flow_gain,
        "update_block.mask.2.weight": lambda t: t * 100.0,
        "update_block.mask.2.bias": lambda t: t * 100.0,
    })
    d = torch.arange(160).view(1, -1) - torch.arange(160).view(-1, 1)
    sd["att.pos_emb.rel_ind"] = d + 159
    return sd


def motion_encoder(flow, corr23, corr21, sd, p="update_block.encoder"):
    """SKMotionEncoder6_Deep_nopool_res.forward, sk2.py:46-58 (corr = cat([corr23, corr21]))."""
    k = (1, 15)
    c1 = omof.pcblock(corr23, sd, p + ".convc1", k)
    c2 = omof.pcblock(corr21, sd, p + ".convc1", k)
    cor = omof.pcblock(F.gelu(torch.cat([c1, c2], 1)), sd, p + ".convc2", k)
    flo = omof.pcblock(omof._conv(flow, sd, p + ".convf1_"), sd, p + ".convf2", k)
    out = omof.pcblock(torch.cat([cor, flo], 1), sd, p + ".conv", k)
    return torch.cat([out, flow], 1)


def update_block(net, inp, corr23, corr21, flow, attn, sd, with_mask=True):
    """SKUpdateBlock6_Deep_nopoolres_AllDecoder2.forward, sk2.py:80-92."""
    p = "update_block"
    mf = motion_encoder(flow, corr23, corr21, sd)
    mfg = omof.gma_aggregate(attn, mf, sd)
    net = omof.pcblock(torch.cat([net, inp, mf, mfg], 1), sd, p + ".gru", (1, 7))
    delta = omof.pcblock(net, sd, p + ".flow_head", (1, 15))
    mask = None
    if with_mask:
        m = F.relu(omof._conv(net, sd, p + ".mask.0", padding=1))
        mask = 0.25 * omof._conv(m, sd, p + ".mask.2")
    return net, mask, delta


def bof_encode(images, sd, encoder=None):
    """images [1,3,3,H,W] (raw caller values) -> fmaps [3,256,h,w], net/inp [1,128,h,w]. network.py:86-112"""
    return omof.mof_encode(images, sd, encoder)       # T = 3: cnet sees frame 1 only, as network.py:108


def bof_refine(fmaps, net, inp, sd, iters=ITERS):
    """-> [1, 2, 2, 8h, 8w]: stack([flow_up_23, flow_up_21], dim=1).  network.py:95-145 with B = 1."""
    h, w = fmaps.shape[-2:]
    pyr21 = omof.corr_pyramid(fmaps[1:2], fmaps[0:1])
    pyr23 = omof.corr_pyramid(fmaps[1:2], fmaps[2:3])
    attn = omof.gma_attention(inp, sd)
    c0 = omof.pixel_grid(1, h, w, fmaps.device)
    c21, c23 = c0.clone(), c0.clone()
    out = None
    for it in range(iters):
        corr21 = omof.corr_lookup(pyr21, c21)
        corr23 = omof.corr_lookup(pyr23, c23)
        flow = torch.cat([c23 - c0, c21 - c0], 1)
        last = it == iters - 1
        net, mask, delta = update_block(net, inp, corr23, corr21, flow, attn, sd, with_mask=last)
        c23 = c23 + delta[:, 0:2]
        c21 = c21 + delta[:, 2:4]
        if last:
            m21, m23 = mask[:, :576], mask[:, 576:]                 # network.py:130
            out = torch.stack([omof.convex_upsample8(c23 - c0, m23), omof.convex_upsample8(c21 - c0, m21)], 1)
    return out


def bof_forward(images, sd, iters=ITERS, encoder=None):
    fmaps, net, inp = bof_encode(images, sd, encoder)
    return bof_refine(fmaps, net, inp, sd, iters)


def core_compute_flow(window, sd, iters=ITERS, encoder=None):
    """What VideoFlowCore.compute_flow_from_tensor returns for a BOF model: pad, forward, unpad, index
    shape[1] // 2 = 1, i.e. the 2->1 (backward) flow.  processing/videoflow_core.py:181-198"""
    H, W = window.shape[-2:]
    flows = omof.unpad(bof_forward(omof.pad_frames(window), sd, iters, encoder), H, W)
    return flows[0, flows.shape[1] // 2]
