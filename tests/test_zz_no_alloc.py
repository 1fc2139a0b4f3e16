"""vfml_corr_volume_pyramid_tiled_static: the allocation-free variant of the tiled volume + pyramid build (scratch in a
device array built into libvfml.so) against the existing entry point with a caller workspace, both through the C ABI.
Collected last on purpose (the file name)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _levels(lib, n, h, w):
    return [torch.full((n * h * w, lib.vfml_corr_tiled_slab_bytes(h, w, l)), 0xAB, dtype=torch.uint8, device="cuda") for l in range(4)]


def _ptrs(levels):
    return (C.c_void_p * 4)(*[l.data_ptr() for l in levels])


def _both(lib, check, n, Cc, h, w, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    f1 = torch.randn((n, Cc, h, w), generator=g).cuda()
    f2 = torch.randn((n, Cc, h, w), generator=g).cuda()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    need = lib.vfml_corr_volume_tiled_workspace(n, Cc, h, w)
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    want, got = _levels(lib, n, h, w), _levels(lib, n, h, w)
    check(lib.vfml_corr_volume_pyramid_tiled(f1.data_ptr(), f2.data_ptr(), n, Cc, h, w, _ptrs(want), 4, ws.data_ptr(), need, stream),
          "corr_volume_pyramid_tiled")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    check(lib.vfml_corr_volume_pyramid_tiled_static(f1.data_ptr(), f2.data_ptr(), n, Cc, h, w, _ptrs(got), 4, stream),
          "corr_volume_pyramid_tiled_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return want, got, free0, free1


def test_static_scratch_matches_caller_workspace_twice():
    from video_flow_ml_b200 import _lib
    lib = _lib.load()
    assert lib.vfml_corr_volume_pyramid_tiled_static_max_bytes() == 16 << 20
    # two calls in a row on the shared array: a larger shape first, then a smaller one over its leftovers, then the first again
    free = []
    for n, Cc, h, w, seed in ((1, 256, 40, 64, 1), (2, 128, 24, 40, 2), (1, 256, 40, 64, 3)):
        want, got, f0, f1 = _both(lib, _lib.check, n, Cc, h, w, seed)
        for l in range(4):
            assert torch.equal(got[l], want[l]), f"level {l} differs for shape {(n, Cc, h, w)}"     # same kernels: bit-exact
        assert int((want[0] != 0xAB).sum()) > 0
        free.append((f0, f1))
    # device memory in use does not grow across the later calls (the first may load kernels lazily)
    for f0, f1 in free[1:]:
        assert f1 >= f0, f"free device memory fell from {f0} to {f1} across an allocation-free call"


def test_static_scratch_rejects_shapes_beyond_the_array():
    from video_flow_ml_b200 import _lib
    lib = _lib.load()
    n, Cc, h, w = 1, 256, 55, 128                       # the C1 tile: 100 MB of level-0 volume
    assert lib.vfml_corr_volume_tiled_workspace(n, Cc, h, w) > lib.vfml_corr_volume_pyramid_tiled_static_max_bytes()
    f = torch.zeros((n, Cc, h, w), device="cuda")
    lv = _levels(lib, n, h, w)
    rc = lib.vfml_corr_volume_pyramid_tiled_static(f.data_ptr(), f.data_ptr(), n, Cc, h, w, _ptrs(lv), 4,
                                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -4                                     # VFML_ERR_WORKSPACE
    msg = lib.vfml_last_error().decode()
    assert "scratch" in msg and "16777216" in msg
    torch.cuda.synchronize()
    assert all(int((l != 0xAB).sum()) == 0 for l in lv)  # nothing was launched
