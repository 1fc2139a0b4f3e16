"""Drop-in ``CorrBlock`` (VideoFlow/core/Networks/MOFNetStack/corr.py:59-107) on libvfml kernels:
same constructor, ``corr_pyramid`` attribute, ``__call__`` and static ``corr``."""
import torch

from . import ops


class CorrBlock:
    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels = num_levels
        self.radius = radius
        # list of [B*h*w, 1, h_l, w_l] fp32, each level avg_pool2d(2, 2) of the previous (corr.py:60-74)
        self.corr_pyramid = ops.corr_volume_pyramid(fmap1.float().contiguous(), fmap2.float().contiguous(), num_levels)

    def __call__(self, coords):
        """coords [B, 2, h, w] -> [B, num_levels*(2r+1)^2, h, w] fp32 contiguous (corr.py:76-97, tap order Q13)."""
        return ops.corr_lookup(self.corr_pyramid, coords.float().contiguous(), self.radius)

    @staticmethod
    def corr(fmap1, fmap2):
        """All-pairs correlation [B, h, w, 1, h, w] = fmap1^T fmap2 / sqrt(C) (corr.py:99-107)."""
        b, _, h, w = fmap1.shape
        vol = ops.corr_volume_pyramid(fmap1.float().contiguous(), fmap2.float().contiguous(), 1)[0]
        return vol.view(b, h, w, 1, h, w)


class AlternateCorrBlock:
    """Volume-free lookup block (MOFNetStack/corr.py:149-177): same constructor and ``__call__`` as the
    reference; ``alt_cuda_corr.forward`` is ``ops.alt_corr_forward`` (vfml_alt_corr_forward)."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels = num_levels
        self.radius = radius
        self.pyramid = [(fmap1, fmap2)]
        for _ in range(self.num_levels):                     # corr.py:155-158: only fmap2 is pooled
            fmap2 = torch.nn.functional.avg_pool2d(fmap2, 2, stride=2)
            self.pyramid.append((None, fmap2))
        # NHWC copies are made once here instead of once per call (corr.py:168-169)
        self._f1 = fmap1.permute(0, 2, 3, 1).contiguous()
        self._f2 = [self.pyramid[i][1].permute(0, 2, 3, 1).contiguous() for i in range(self.num_levels)]

    def __call__(self, coords):
        """coords [B, 2, h, w] -> [B, num_levels*(2r+1)^2, h, w] fp32 (corr.py:160-177)."""
        coords = coords.permute(0, 2, 3, 1)
        B, H, W, _ = coords.shape
        dim = self.pyramid[0][0].shape[1]
        corr_list = []
        for i in range(self.num_levels):
            coords_i = (coords.float() / 2 ** i).reshape(B, 1, H, W, 2).contiguous()
            corr, = ops.alt_corr_forward(self._f1, self._f2[i], coords_i, self.radius)
            corr_list.append(corr.squeeze(1))
        corr = torch.stack(corr_list, dim=1).reshape(B, -1, H, W)
        return corr / torch.sqrt(torch.tensor(dim).float())
