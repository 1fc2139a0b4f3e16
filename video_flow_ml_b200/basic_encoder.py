"""MemFlow's BasicEncoder (MemFlow/core/Networks/MemFlowNet/cnn.py:5-56,117-203), eval mode, on torch library ops.

Boundary input of the MemFlow hot path, like the Twins stages for VideoFlow: the 7x7 / 3x3
strided convolutions and the norms are library calls; everything after the encoders is libvfml.
``norm='instance'`` (fnet): InstanceNorm2d without affine parameters; ``norm='batch'`` (cnet): BatchNorm2d with
running statistics, folded into a per-channel scale and shift at load time.
"""
import torch
import torch.nn.functional as F


class BasicEncoder:
    def __init__(self, sd, prefix, device, norm, dtype=torch.float32):
        if norm not in ("instance", "batch"):
            raise ValueError("norm must be 'instance' or 'batch'")
        self.norm, self.dt, self.p = norm, dtype, prefix
        g = lambda k: sd[f"{prefix}.{k}"].to(device=device, dtype=torch.float32)
        self.w = {}
        for k in sd:
            if k.startswith(prefix + ".") and (k.endswith("weight") or k.endswith("bias")) and ".norm" not in k and ".downsample.1" not in k:
                self.w[k[len(prefix) + 1:]] = sd[k].to(device=device, dtype=dtype)
        self.bn = {}
        if norm == "batch":
            names = {k[len(prefix) + 1:-len(".running_mean")] for k in sd if k.startswith(prefix + ".") and k.endswith(".running_mean")}
            for n in names:
                scale = g(n + ".weight") / torch.sqrt(g(n + ".running_var") + 1e-5)
                shift = g(n + ".bias") - g(n + ".running_mean") * scale
                self.bn[n] = (scale.view(1, -1, 1, 1).to(dtype), shift.view(1, -1, 1, 1).to(dtype))

    def _norm(self, x, name):
        if self.norm == "batch":
            s, b = self.bn[name]
            return x * s + b
        return F.instance_norm(x, eps=1e-5)

    def _conv(self, x, name, **kw):
        return F.conv2d(x, self.w[name + ".weight"], self.w[name + ".bias"], **kw)

    def _block(self, x, q, stride):
        y = F.relu(self._norm(self._conv(x, q + ".conv1", stride=stride, padding=1), q + ".norm1"))
        y = F.relu(self._norm(self._conv(y, q + ".conv2", padding=1), q + ".norm2"))
        if stride != 1:
            x = self._norm(self._conv(x, q + ".downsample.0", stride=stride), q + ".norm3")
        return F.relu(x + y)

    @torch.no_grad()
    def __call__(self, x):
        """[N, 3, H, W] -> [N, 256, H/8, W/8] fp32"""
        if self.dt != torch.float32 or not x.is_cuda:
            return self._forward(x)
        # fp32 tier: true fp32 library math (cuDNN defaults to TF32 for convolutions)
        c = torch.backends.cudnn.allow_tf32
        torch.backends.cudnn.allow_tf32 = False
        try:
            return self._forward(x)
        finally:
            torch.backends.cudnn.allow_tf32 = c

    def _forward(self, x):
        x = F.relu(self._norm(self._conv(x.to(self.dt), "conv1", stride=2, padding=3), "norm1"))
        for li, stride in ((1, 1), (2, 2), (3, 2)):
            x = self._block(x, f"layer{li}.0", stride)
            x = self._block(x, f"layer{li}.1", 1)
        return self._conv(x, "conv2").float()
