"""Truncated Twins-SVT-Large feature/context encoder as PyTorch library ops (boundary input of
the hot path).

The reference builds it with ``timm.create_model('twins_svt_large')`` and keeps stages 0-1
(VideoFlow/core/Networks/encoders.py:7-40); timm is a third-party dependency that is not in
the reference tree, so the math follows the class definitions vendored in
VideoFlow/core/Networks/twins_ft.py: PatchEmbed :361-386, LocallyGroupedAttn :214-261,
GlobalSubSampleAttn :264-312, Block :315-338, PosConv :341-355, svt_large :509-515.
Weights keep timm's key names (``<prefix>.patch_embeds.0.proj.weight`` ...), so upstream
checkpoints load unchanged.  LayerNorm, the PEG depthwise convolution and (fp16 tier) both attention types are libvfml
kernels; the linear layers and the strided patch / sub-sampling convolutions are library calls (cuBLAS / cuDNN).  The
outputs are per-image, so VideoFlowCore.compute_flow_from_frames caches them per frame across sliding windows.
"""
import torch
import torch.nn.functional as F

_DIMS = (128, 256)
_HEADS = (4, 8)
_SR = (8, 4)
_PATCH = (4, 2)
_WS = 7


def _ops():
    from . import ops
    return ops


class TwinsStage01:
    def __init__(self, state_dict, prefix, device, dtype=torch.float32):
        n = len(prefix) + 1
        self.p = {k[n:]: v.to(device=device, dtype=dtype) for k, v in state_dict.items()
                  if k.startswith(prefix + ".") and v.is_floating_point()}
        self.dtype = dtype
        # PEG weights tap-major [9, C] for the libvfml kernel
        self.peg = {s_: self.p[f"pos_block.{s_}.proj.0.weight"].reshape(-1, 9).t().contiguous() for s_ in range(2)
                    if f"pos_block.{s_}.proj.0.weight" in self.p}

    def _ln(self, x, name, eps):
        w, b = self.p[name + ".weight"], self.p[name + ".bias"]
        if x.is_cuda and x.dtype == w.dtype and x.shape[-1] % 8 == 0 and x.shape[-1] <= 512:
            return _ops().layer_norm_rows(x.contiguous(), w, b, eps)      # libvfml kernel; strided views are packed first, as torch does
        return F.layer_norm(x, (x.shape[-1],), w, b, eps)

    def _fc(self, x, name):
        return F.linear(x, self.p[name + ".weight"], self.p.get(name + ".bias"))

    def _lsa(self, x, H, W, name, heads):
        """Locally-grouped self-attention over 7x7 windows, zero padded at right/bottom."""
        B, N, C = x.shape
        x = x.view(B, H, W, C)
        pr, pb = (-W) % _WS, (-H) % _WS
        if pr or pb:
            x = F.pad(x, (0, 0, 0, pr, 0, pb))
        gh, gw = (H + pb) // _WS, (W + pr) // _WS
        x = x.view(B, gh, _WS, gw, _WS, C).permute(0, 1, 3, 2, 4, 5).reshape(B * gh * gw, _WS * _WS, C)
        qkv = self._fc(x, name + ".qkv")                                        # [windows, 49, 3C]
        if qkv.is_cuda and qkv.dtype == torch.float16 and C // heads == 32:      # libvfml fused attention on the q / k / v slices in place
            o = _ops().attention_hd32(qkv[..., 0:C], qkv[..., C:2 * C], qkv[..., 2 * C:3 * C], heads)
        else:
            qkv = qkv.view(-1, _WS * _WS, 3, heads, C // heads).permute(2, 0, 3, 1, 4)
            o = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2]).transpose(1, 2)
        o = o.reshape(B, gh, gw, _WS, _WS, C).permute(0, 1, 3, 2, 4, 5)
        o = o.reshape(B, gh * _WS, gw * _WS, C)[:, :H, :W].reshape(B, N, C)
        return self._fc(o, name + ".proj")

    def _gsa(self, x, H, W, name, heads, sr):
        """Global attention against keys/values sub-sampled by a strided conv."""
        B, N, C = x.shape
        hd = C // heads
        q = self._fc(x, name + ".q")
        # .contiguous(): a [1, C, H, W] view of token-major data is stride-ambiguous (NCHW or channels-last); the library's
        # kernel choice for such views has produced wrong results at batch 1 (see ops.peg3x3)
        y = F.conv2d(x.transpose(1, 2).reshape(B, C, H, W).contiguous(), self.p[name + ".sr.weight"], self.p[name + ".sr.bias"], stride=sr)
        y = self._ln(y.flatten(2).transpose(1, 2), name + ".norm", 1e-5)
        kv = self._fc(y, name + ".kv")                                           # [B, Nk, 2C]
        if kv.is_cuda and kv.dtype == torch.float16 and hd == 32:
            o = _ops().attention_hd32(q, kv[..., 0:C], kv[..., C:2 * C], heads)
        else:
            kv = kv.view(B, -1, 2, heads, hd).permute(2, 0, 3, 1, 4)
            o = F.scaled_dot_product_attention(q.view(B, N, heads, hd).transpose(1, 2), kv[0], kv[1]).transpose(1, 2).reshape(B, N, C)
        return self._fc(o, name + ".proj")

    @torch.no_grad()
    def __call__(self, x):
        """x [B, 3, H, W] in [-1, 1] -> [B, 256, H/8, W/8] fp32."""
        if self.dtype != torch.float32 or not x.is_cuda:
            return self._forward(x)
        # fp32 tier: true fp32 library math (cuDNN / cuBLAS default to TF32 for convolutions)
        c, m = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        try:
            return self._forward(x)
        finally:
            torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = c, m

    def _forward(self, x):
        x = x.to(self.dtype)
        B = x.shape[0]
        for s in range(2):
            pe = f"patch_embeds.{s}"
            x = F.conv2d(x, self.p[pe + ".proj.weight"], self.p[pe + ".proj.bias"], stride=_PATCH[s])
            H, W = x.shape[-2:]
            x = self._ln(x.flatten(2).transpose(1, 2), pe + ".norm", 1e-5)
            for j in range(2):
                blk = f"blocks.{s}.{j}"
                y = self._ln(x, blk + ".norm1", 1e-6)
                if j == 0:
                    x = x + self._lsa(y, H, W, blk + ".attn", _HEADS[s])
                else:
                    x = x + self._gsa(y, H, W, blk + ".attn", _HEADS[s], _SR[s])
                y = self._ln(x, blk + ".norm2", 1e-6)
                x = x + self._fc(F.gelu(self._fc(y, blk + ".mlp.fc1")), blk + ".mlp.fc2")
                if j == 0:   # PEG after the first block of each stage (encoders.py:27-29, :35-37)
                    pb = f"pos_block.{s}.proj.0"
                    if x.is_cuda:      # libvfml kernel on the token-major layout (no NCHW round trip)
                        x = _ops().peg3x3(x.contiguous(), B, H, W, self.peg[s], self.p[pb + ".bias"])
                    else:
                        m = x.transpose(1, 2).reshape(B, -1, H, W)
                        m = m + F.conv2d(m, self.p[pb + ".weight"], self.p[pb + ".bias"], padding=1, groups=m.shape[1])
                        x = m.flatten(2).transpose(1, 2)
            x = x.transpose(1, 2).reshape(B, -1, H, W).contiguous()
        return x.float().contiguous()
