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
