"""CPU restatement of the truncated Twins-SVT-Large encoder (oracle; test infrastructure only).

The reference builds its feature/context encoders with
``timm.create_model('twins_svt_large')`` (VideoFlow/core/Networks/encoders.py:7-19)
and runs only stages 0 and 1 (encoders.py:21-40).  ``timm==0.4.12`` is a
third-party dependency that is NOT in the reference tree and cannot be
installed offline, so the math is restated here from the vendored class
definitions the reference keeps in ``VideoFlow/core/Networks/twins_ft.py``:

* PatchEmbed            twins_ft.py:361-386 (conv k=s=patch, flatten, LayerNorm eps 1e-5)
* LocallyGroupedAttn    twins_ft.py:214-261 (7x7 windows, zero padding right/bottom)
* GlobalSubSampleAttn   twins_ft.py:264-312 (sr conv + LayerNorm eps 1e-5 on keys/values)
* Block                 twins_ft.py:315-338 (pre-norm, LayerNorm eps 1e-6: twins_ft.py:398)
* PosConv (PEG)         twins_ft.py:341-355 (depthwise 3x3 + residual)
* svt_large hyper-parameters twins_ft.py:509-515

composed exactly as encoders.py:21-40 (PEG after block 0 of each stage, stages
0-1 only, token->NCHW reshape after each stage).  PARITY UNPINNED at this
boundary: no reference test or golden vector pins the encoder output.

Weights are a flat dict with timm's key names under a prefix, e.g.
``fnet.svt.patch_embeds.0.proj.weight``.
"""
import torch
import torch.nn.functional as F

EMBED_DIMS = (128, 256)
NUM_HEADS = (4, 8)
SR_RATIOS = (8, 4)
WINDOW = 7
PATCH = (4, 2)


def _ln(x, sd, p, eps):
    return F.layer_norm(x, (x.shape[-1],), sd[p + ".weight"], sd[p + ".bias"], eps)


def _lin(x, sd, p):
    return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))


def _window_attention(x, hw, sd, p, heads):
    # twins_ft.py:233-261
    B, N, C = x.shape
    H, W = hw
    ws = WINDOW
    x = x.view(B, H, W, C)
    pr = (ws - W % ws) % ws
    pb = (ws - H % ws) % ws
    x = F.pad(x, (0, 0, 0, pr, 0, pb))
    Hp, Wp = H + pb, W + pr
    gh, gw = Hp // ws, Wp // ws
    x = x.reshape(B, gh, ws, gw, ws, C).transpose(2, 3)          # B gh gw ws ws C
    qkv = _lin(x, sd, p + ".qkv").reshape(B, gh * gw, ws * ws, 3, heads, C // heads)
    qkv = qkv.permute(3, 0, 1, 4, 2, 5)
    q, k, v = qkv[0], qkv[1], qkv[2]
    a = (q @ k.transpose(-2, -1)) * ((C // heads) ** -0.5)
    a = a.softmax(dim=-1)
    o = (a @ v).transpose(2, 3).reshape(B, gh, gw, ws, ws, C)
    o = o.transpose(2, 3).reshape(B, gh * ws, gw * ws, C)
    if pr or pb:
        o = o[:, :H, :W, :].contiguous()
    return _lin(o.reshape(B, N, C), sd, p + ".proj")


def _subsampled_attention(x, hw, sd, p, heads, sr):
    # twins_ft.py:291-312
    B, N, C = x.shape
    hd = C // heads
    q = _lin(x, sd, p + ".q").reshape(B, N, heads, hd).permute(0, 2, 1, 3)
    y = x.permute(0, 2, 1).reshape(B, C, *hw)
    y = F.conv2d(y, sd[p + ".sr.weight"], sd[p + ".sr.bias"], stride=sr)
    y = y.reshape(B, C, -1).permute(0, 2, 1)
    y = _ln(y, sd, p + ".norm", 1e-5)
    kv = _lin(y, sd, p + ".kv").reshape(B, -1, 2, heads, hd).permute(2, 0, 3, 1, 4)
    k, v = kv[0], kv[1]
    a = ((q @ k.transpose(-2, -1)) * (hd ** -0.5)).softmax(dim=-1)
    o = (a @ v).transpose(1, 2).reshape(B, N, C)
    return _lin(o, sd, p + ".proj")


def _block(x, hw, sd, p, heads, sr, local):
    # twins_ft.py:335-338
    y = _ln(x, sd, p + ".norm1", 1e-6)
    if local:
        x = x + _window_attention(y, hw, sd, p + ".attn", heads)
    else:
        x = x + _subsampled_attention(y, hw, sd, p + ".attn", heads, sr)
    y = _ln(x, sd, p + ".norm2", 1e-6)
    y = _lin(F.gelu(_lin(y, sd, p + ".mlp.fc1")), sd, p + ".mlp.fc2")
    return x + y


def _peg(x, hw, sd, p):
    # twins_ft.py:348-355
    B, N, C = x.shape
    m = x.transpose(1, 2).reshape(B, C, *hw)
    m = F.conv2d(m, sd[p + ".proj.0.weight"], sd[p + ".proj.0.bias"], padding=1, groups=C) + m
    return m.flatten(2).transpose(1, 2)


def twins_svt_forward(x, sd, prefix):
    """images [B,3,H,W] (already scaled to [-1,1]) -> features [B,256,H/8,W/8].

    Follows encoders.py:21-40 with layer=2.  ``prefix`` is e.g. ``"fnet.svt"``.
    """
    B = x.shape[0]
    for s in range(2):
        p = PATCH[s]
        x = F.conv2d(x, sd[f"{prefix}.patch_embeds.{s}.proj.weight"],
                     sd[f"{prefix}.patch_embeds.{s}.proj.bias"], stride=p)
        hw = (x.shape[2], x.shape[3])
        x = x.flatten(2).transpose(1, 2)
        x = _ln(x, sd, f"{prefix}.patch_embeds.{s}.norm", 1e-5)
        for j in range(2):
            x = _block(x, hw, sd, f"{prefix}.blocks.{s}.{j}", NUM_HEADS[s], SR_RATIOS[s], local=(j == 0))
            if j == 0:
                x = _peg(x, hw, sd, f"{prefix}.pos_block.{s}")
        x = x.reshape(B, *hw, -1).permute(0, 3, 1, 2).contiguous()
    return x


def twins_param_shapes(prefix):
    """Key -> shape of every parameter timm's truncated twins_svt_large exposes
    (after the ``del`` statements of encoders.py:12-19; ``svt.norm`` of the
    deleted last stage survives)."""
    out = {}
    cin = 3
    for s in range(2):
        C = EMBED_DIMS[s]
        p = PATCH[s]
        out[f"{prefix}.patch_embeds.{s}.proj.weight"] = (C, cin, p, p)
        out[f"{prefix}.patch_embeds.{s}.proj.bias"] = (C,)
        out[f"{prefix}.patch_embeds.{s}.norm.weight"] = (C,)
        out[f"{prefix}.patch_embeds.{s}.norm.bias"] = (C,)
        for j in range(2):
            b = f"{prefix}.blocks.{s}.{j}"
            for n in ("norm1", "norm2"):
                out[f"{b}.{n}.weight"] = (C,)
                out[f"{b}.{n}.bias"] = (C,)
            if j == 0:
                out[f"{b}.attn.qkv.weight"] = (3 * C, C)
                out[f"{b}.attn.qkv.bias"] = (3 * C,)
            else:
                out[f"{b}.attn.q.weight"] = (C, C)
                out[f"{b}.attn.q.bias"] = (C,)
                out[f"{b}.attn.kv.weight"] = (2 * C, C)
                out[f"{b}.attn.kv.bias"] = (2 * C,)
                sr = SR_RATIOS[s]
                out[f"{b}.attn.sr.weight"] = (C, C, sr, sr)
                out[f"{b}.attn.sr.bias"] = (C,)
                out[f"{b}.attn.norm.weight"] = (C,)
                out[f"{b}.attn.norm.bias"] = (C,)
            out[f"{b}.attn.proj.weight"] = (C, C)
            out[f"{b}.attn.proj.bias"] = (C,)
            out[f"{b}.mlp.fc1.weight"] = (4 * C, C)
            out[f"{b}.mlp.fc1.bias"] = (4 * C,)
            out[f"{b}.mlp.fc2.weight"] = (C, 4 * C)
            out[f"{b}.mlp.fc2.bias"] = (C,)
        out[f"{prefix}.pos_block.{s}.proj.0.weight"] = (C, 1, 3, 3)
        out[f"{prefix}.pos_block.{s}.proj.0.bias"] = (C,)
        cin = C
    out[f"{prefix}.norm.weight"] = (1024,)
    out[f"{prefix}.norm.bias"] = (1024,)
    return out
