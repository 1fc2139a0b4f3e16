"""CPU stand-in for ``video_flow_ml_b200.ops`` (TEST DOUBLE, tests only).

Implements the same call signatures with plain torch CPU ops so that the host-side
orchestration of the engine (buffer carving, column ranges, call order, weight prepacking) can
be checked against the oracle in the CPU test suite.  It is never importable from the product
package; the product always binds libvfml.so and has no CPU path.
"""
import torch
import torch.nn.functional as F


def nchw_to_rows(src, dst):
    n, c, h, w = src.shape
    dst.copy_(src.permute(0, 2, 3, 1).reshape(n * h * w, c).to(dst.dtype))
    return dst


def rows_to_nchw(src, dst):
    n, c, h, w = dst.shape
    dst.copy_(src.float().reshape(n, h, w, c).permute(0, 3, 1, 2))
    return dst


def corr_volume_pyramid(f1, f2, num_levels=4, vol_dtype=torch.float32):
    n, C, h, w = f1.shape
    v = torch.matmul(f1.reshape(n, C, h * w).transpose(1, 2), f2.reshape(n, C, h * w)) / C ** 0.5
    v = v.reshape(n * h * w, 1, h, w)
    out = [v.to(vol_dtype)]
    for _ in range(num_levels - 1):
        v = F.avg_pool2d(v, 2, stride=2)
        out.append(v.to(vol_dtype))
    return out


def corr_lookup(levels, coords, radius=4, out=None, out_dtype=torch.float32):
    n, _, h, w = coords.shape
    r = radius
    d = torch.arange(-r, r + 1, dtype=torch.float32)
    res = []
    for l, vol in enumerate(levels):
        vol = vol.float()
        H, W = vol.shape[-2:]
        c = coords.permute(0, 2, 3, 1).reshape(n * h * w, 1, 1, 2) / 2 ** l
        gx = c[..., 0] + d.view(1, -1, 1)      # first (slow) tap index moves x
        gy = c[..., 1] + d.view(1, 1, -1)
        gx, gy = torch.broadcast_tensors(gx, gy)
        grid = torch.stack([2 * gx / (W - 1) - 1, 2 * gy / (H - 1) - 1], -1)
        res.append(F.grid_sample(vol, grid, align_corners=True).view(n * h * w, -1))
    rows = torch.cat(res, 1)
    if out is None:
        return rows.view(n, h, w, -1).permute(0, 3, 1, 2).contiguous().to(out_dtype)
    out.copy_(rows.to(out.dtype))
    return out


def corr_lookup_pair(levels_f, levels_b, coords_f, coords_b, out_f, out_b):
    corr_lookup(levels_f, coords_f, 4, out=out_f)
    corr_lookup(levels_b, coords_b, 4, out=out_b)


def corr_lookup_perm(device=None):
    from video_flow_ml_b200.ops import corr_lookup_perm as real      # pure index arithmetic, no library call
    return real(device)


class TiledPyramid:
    """Stand-in for ops.TiledPyramid: levels[l] is [n*h*w, 1, h_l, w_l] (row = source pixel), so slots of source frames are
    row ranges like in the real layout."""

    def __init__(self, levels, n, h, w):
        self.levels, self.n, self.h, self.w = levels, n, h, w


def tiled_pyramid_alloc(n, h, w, device):
    return TiledPyramid([torch.full((n * h * w, 1, h >> l, w >> l), float("nan")) for l in range(4)], n, h, w)


def tiled_volume_workspace(n, Cc, h, w, device):
    return None


def corr_volume_pyramid_tiled(f1, f2, workspace=None, out=None):
    levels = corr_volume_pyramid(f1, f2, 4)
    if out is None:
        return TiledPyramid(levels, f1.shape[0], f1.shape[2], f1.shape[3])
    for dst, src in zip(out.levels, levels):
        dst.copy_(src)
    return out


def corr_lookup_tiled(pyr_a, pyr_b, coords_a, coords_b, out_a, out_b, math_f32=False):
    """Same taps as corr_lookup, written in the tiled kernel's permuted channel order."""
    perm = corr_lookup_perm()
    for pyr, c, out in ((pyr_a, coords_a, out_a), (pyr_b, coords_b, out_b)):
        if pyr is None:
            continue
        tmp = torch.empty(out.shape, dtype=torch.float32)
        corr_lookup(pyr.levels, c, 4, out=tmp)
        out.copy_(tmp[:, perm].to(out.dtype))
    return out_a, out_b


def convex_upsample8(flow, mask, out=None, crop=None, flow_is_coords=False):
    n, _, h, w = flow.shape
    if flow_is_coords:
        ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
        flow = flow - torch.stack([xs, ys], 0).float()[None]
    if mask.dim() == 2:
        mask = mask.float().reshape(n, h, w, 576).permute(0, 3, 1, 2)
    m = torch.softmax(mask.reshape(n, 1, 9, 8, 8, h, w), dim=2)
    u = F.unfold(8 * flow, [3, 3], padding=1).view(n, 2, 9, 1, 1, h, w)
    u = torch.sum(m * u, dim=2).permute(0, 1, 4, 2, 5, 3).reshape(n, 2, 8 * h, 8 * w)
    if crop is not None:
        t, l, oh, ow = crop
        u = u[:, :, t:t + oh, l:l + ow]
    if out is None:
        return u.contiguous()
    out.copy_(u)
    return out


def gemm(a, w, out, bias=None, res=None, act=None, dw=None, alpha=1.0, addto=False, require_tc=False, weights=False):
    v = alpha * (a.float() @ w.float().t())
    if bias is not None:
        v = v + bias
    if res is not None:
        v = v + res.float()
    if act == "gelu":
        v = F.gelu(v)
    elif act == "relu":
        v = F.relu(v)
    if dw is not None:
        v = F.gelu(v + (dw[0] * v + dw[1]))
    if addto:
        v = v + out.float()
    out.copy_(v.to(out.dtype))
    return out


def dwconv_res_gelu(x, y, n_img, h, w, weight_tapmajor, bias, k):
    C = x.shape[1]
    xi = x.float().reshape(n_img, h, w, C).permute(0, 3, 1, 2)
    wt = weight_tapmajor.t().reshape(C, 1, k, k)
    v = F.gelu(xi + F.conv2d(xi, wt, bias, padding=k // 2, groups=C))
    y.copy_(v.permute(0, 2, 3, 1).reshape(n_img * h * w, C).to(y.dtype))
    return y


class gemm_autotuning:
    """no-op stand-in for ops.gemm_autotuning (kernel selection has no CPU counterpart)"""

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def gemm_splitk(a, w, out, n_splits, partial, res=None, alpha=1.0):
    Ks = a.shape[1] // n_splits
    acc = torch.zeros(out.shape, dtype=torch.float32)
    for s_ in range(n_splits):
        acc += alpha * (a[:, s_ * Ks:(s_ + 1) * Ks].float() @ w[:, s_ * Ks:(s_ + 1) * Ks].float().t())
    if res is not None:
        acc += res.float()
    out.copy_(acc.to(out.dtype))
    return out


def softmax_rows(src, dst):
    dst.copy_(torch.softmax(src, dim=-1).to(dst.dtype))
    return dst


def _sample(img, cx, cy):
    H, W = img.shape[-2:]
    grid = torch.stack([2 * cx / (W - 1) - 1, 2 * cy / (H - 1) - 1], -1)
    return F.grid_sample(img, grid, align_corners=True)


def motion_glue(fc, bc, x400, f128, mf, w1, b1, n_frames, h, w, rot=0):
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    g = torch.stack([xs, ys], 0).float()[None]
    flow4 = torch.cat([fc - g, bc - g], 1)                                     # [F,4,h,w]
    mhs = x400[:, 352:400].float().reshape(n_frames, h, w, 48).permute(0, 3, 1, 2)
    z = torch.zeros(48, h, w)
    # ring order: slot f holds the window's frame (f - rot) mod F; its neighbours in window order are the adjacent slots
    logical = [(f - rot) % n_frames for f in range(n_frames)]
    nxt = torch.stack([mhs[(f + 1) % n_frames] if logical[f] + 1 < n_frames else z for f in range(n_frames)])
    prv = torch.stack([mhs[(f - 1) % n_frames] if logical[f] > 0 else z for f in range(n_frames)])
    fw = _sample(nxt, fc[:, 0], fc[:, 1])
    bw = _sample(prv, bc[:, 0], bc[:, 1])
    rows = lambda t: t.permute(0, 2, 3, 1).reshape(n_frames * h * w, -1)
    x400[:, 256:304] = rows(fw).to(x400.dtype)
    x400[:, 304:352] = rows(bw).to(x400.dtype)
    f4 = rows(flow4)
    f128.copy_((f4 @ w1.t() + b1).to(f128.dtype))
    mf[:, 124:128] = f4.to(mf.dtype)


def coords_add(fc, bc, delta):
    n, _, h, w = fc.shape
    d = delta.reshape(n, h, w, 4).permute(0, 3, 1, 2)
    fc += d[:, 0:2]
    bc += d[:, 2:4]


def coords_grid(n, h, w, device):
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    return torch.stack([xs, ys], 0).float()[None].repeat(n, 1, 1, 1).contiguous()


def rows_fill(view, values):
    view.copy_(values.to(view.dtype)[None].expand_as(view))


def im2col3x3(x, out, n_img, h, w):
    C = x.shape[1]
    xi = x.float().reshape(n_img, h, w, C).permute(0, 3, 1, 2)
    cols = F.unfold(xi, 3, padding=1).view(n_img, C, 9, h * w)           # [n, C, tap, P]
    out.copy_(cols.permute(0, 3, 2, 1).reshape(n_img * h * w, 9 * C).to(out.dtype))
    return out
