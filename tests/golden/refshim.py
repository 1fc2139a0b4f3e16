"""Import shims that let the UNMODIFIED reference modules run in the build container.

Only used by ``tests/golden/make_golden.py`` (fixture generation) and by the CPU tests that
cross-check the oracle against the live reference when ``/root/reference`` is present.
Nothing here is shipped or timed, and nothing on the GPU box reads ``/root/reference``.

What is shimmed (and why):
* ``timm`` (third-party, absent, pinned ``timm==0.4.12`` in requirements.txt:13).  The shim
  provides the few helper symbols ``VideoFlow/core/Networks/twins_ft.py`` imports, and a
  ``create_model('twins_svt_large')`` that assembles timm's Twins module structure out of
  the reference's OWN vendored class definitions in ``twins_ft.py`` (Block, PosConv,
  PatchEmbed) with the svt_large hyper-parameters of ``twins_ft.py:509-515``.
* ``yacs.config.CfgNode`` (absent): attribute + item dict with ``clone()``.
* ``alt_cuda_corr`` (compiled extension, absent): empty module so ``BOFNet/corr.py:7`` imports.
"""
import importlib
import os
import sys
import types
from functools import partial

import torch
import torch.nn as nn

REF = os.environ.get("VFML_REFERENCE", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF, "VideoFlow", "core"))


class CfgNode(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def clone(self):
        out = CfgNode()
        for k, v in self.items():
            out[k] = v.clone() if isinstance(v, CfgNode) else (list(v) if isinstance(v, list) else v)
        return out


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _Mlp(nn.Module):
    # timm.models.layers.Mlp (0.4.12): fc1 -> act -> drop -> fc2 -> drop
    def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.):
        super().__init__()
        self.fc1 = nn.Linear(in_features, hidden_features or in_features)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_features or in_features, out_features or in_features)
        self.drop = nn.Dropout(drop)

    def forward(self, x):
        return self.drop(self.fc2(self.drop(self.act(self.fc1(x)))))


class _DropPath(nn.Module):
    def __init__(self, p=0.):
        super().__init__()

    def forward(self, x):
        return x


def install():
    if "timm" in sys.modules and getattr(sys.modules["timm"], "_vfml_shim", False):
        return
    _mod("yacs")
    _mod("yacs.config", CfgNode=CfgNode)
    _mod("alt_cuda_corr")
    timm = _mod("timm", _vfml_shim=True)
    timm.__path__ = []
    _mod("timm.data", IMAGENET_DEFAULT_MEAN=(0.485, 0.456, 0.406), IMAGENET_DEFAULT_STD=(0.229, 0.224, 0.225))
    models = _mod("timm.models")
    models.__path__ = []
    _mod("timm.models.layers", Mlp=_Mlp, DropPath=_DropPath,
         to_2tuple=lambda x: x if isinstance(x, tuple) else (x, x),
         trunc_normal_=nn.init.trunc_normal_)
    _mod("timm.models.fx_features", register_notrace_module=lambda c: c)
    _mod("timm.models.registry", register_model=lambda f: f)
    _mod("timm.models.vision_transformer", Attention=object)
    _mod("timm.models.helpers", build_model_with_cfg=None)
    timm.create_model = _create_model
    for p in (os.path.join(REF, "VideoFlow", "core"), os.path.join(REF, "VideoFlow"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)


def _create_model(name, pretrained=False, **kw):
    assert name == "twins_svt_large", name   # pretrained only triggers a download upstream; weights are always replaced by load_state_dict
    tf = importlib.import_module("core.Networks.twins_ft")
    dims, heads, depths, sr, ws = [128, 256, 512, 1024], [4, 8, 16, 32], [2, 2, 18, 2], [8, 4, 2, 1], 7

    class TimmTwins(nn.Module):      # structure of timm 0.4.12 twins.Twins, parts from twins_ft.py
        def __init__(self):
            super().__init__()
            self.depths = depths
            self.patch_embeds = nn.ModuleList()
            self.pos_drops = nn.ModuleList()
            prev, img, patch = 3, (224, 224), 4
            for i in range(4):
                self.patch_embeds.append(tf.PatchEmbed(img, patch, prev, dims[i]))
                self.pos_drops.append(nn.Dropout(0.))
                prev, img, patch = dims[i], tuple(t // patch for t in img), 2
            self.blocks = nn.ModuleList()
            for k in range(4):
                n = depths[k] if k < 2 else 1      # stages 2,3 are deleted by encoders.py:14-17
                self.blocks.append(nn.ModuleList([
                    tf.Block(dim=dims[k], num_heads=heads[k], mlp_ratio=4, sr_ratio=sr[k],
                             norm_layer=partial(nn.LayerNorm, eps=1e-6),
                             ws=1 if i % 2 == 1 else ws) for i in range(n)]))
            self.pos_block = nn.ModuleList([tf.PosConv(d, d) for d in dims])
            self.norm = nn.LayerNorm(dims[-1], eps=1e-6)
            self.head = nn.Linear(dims[-1], 1000)

    return TimmTwins()


def build_mof(iters=32, mixed_precision=False):
    """The reference MOFNet exactly as VideoFlowCore.load_model builds it
    (processing/videoflow_core.py:76,101), with pretrain=False."""
    install()
    cfgmod = importlib.import_module("configs.multiframes_sintel_submission")
    cfg = cfgmod.get_cfg()
    cfg.MOFNetStack.pretrain = False
    cfg.MOFNetStack.mixed_precision = mixed_precision
    cfg.MOFNetStack.decoder_depth = iters
    net = importlib.import_module("core.Networks").build_network(cfg)
    return net.eval(), cfg


def ref_module(name):
    install()
    return importlib.import_module(name)
