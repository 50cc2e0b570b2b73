"""Load the UNMODIFIED reference (/root/reference) on CPU so it can pin the oracle.

Only usable in the authoring container (the GPU box has no /root/reference); it is imported solely by
``tests/golden/make_golden.py`` and by the CPU tests that are skipped when the reference is absent.
Reference files are never copied: they are read where they lie, a handful of hard-coded ``"cuda"``
device strings are substituted in memory at import time, and five missing
third-party modules are stubbed.
"""
import importlib.abc
import importlib.util
import os
import sys
import types

import torch

REF_ROOT = "/root/reference"

_SUBST = {
    "libraries/NeRF/rendering.py": [
        ('device="cuda"', 'device=_DEV()'),
        ("torch.cuda.FloatTensor(batchsize, 1, n, Nf).uniform_()",
         "torch.empty(batchsize, 1, n, Nf, device=_DEV()).uniform_()"),
    ],
    "libraries/NeRF/ray_sampler.py": [('device="cuda"', 'device=_DEV()')],
    "models/generator.py": [(".cuda(homo_img.device)", ".to(homo_img.device)")],
}
_PRELUDE = "import torch as _t; _DEV = lambda: _t.device('cpu'); "


class _Loader(importlib.abc.Loader):
    def __init__(self, path, rel):
        self.path, self.rel = path, rel

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        with open(self.path) as f:
            src = f.read()
        for a, b in _SUBST[self.rel]:
            assert a in src, (self.rel, a)
            src = src.replace(a, b)
        src = _PRELUDE + src  # same line as the file's first statement: line numbers stay stable
        module.__file__ = self.path
        exec(compile(src, self.path, "exec"), module.__dict__)


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        rel = fullname.replace(".", "/") + ".py"
        if rel in _SUBST and os.path.exists(os.path.join(REF_ROOT, rel)):
            return importlib.util.spec_from_loader(fullname, _Loader(os.path.join(REF_ROOT, rel), rel))
        return None


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _EasyDict(dict):
    """Attribute-access dict (stand-in for the missing ``easydict`` package)."""

    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        if isinstance(v, dict) and not isinstance(v, _EasyDict):
            v = _EasyDict(v)
        super().__setitem__(k, v)

    __setattr__ = __setitem__

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)


EasyDict = _EasyDict
_installed = False


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "libraries"))


def install():
    """Make ``import models.generator`` etc. resolve to the reference.  Idempotent."""
    global _installed
    if _installed:
        return
    assert available(), "reference not present"
    os.chdir(REF_ROOT)  # the reference appends relative paths to sys.path
    sys.path.insert(0, REF_ROOT)
    sys.meta_path.insert(0, _Finder())

    ident = type("Identity", (torch.nn.Module,), {"__init__": lambda self, *a, **k: torch.nn.Module.__init__(self),
                                                  "forward": lambda self, x: x})
    k = _stub("kornia")
    k.augmentation = _stub("kornia.augmentation", RandomCrop=ident)
    _stub("pytorch3d")
    names = ["look_at_view_transform", "FoVPerspectiveCameras", "PointLights", "RasterizationSettings",
             "MeshRenderer", "MeshRasterizer", "SoftPhongShader", "Textures", "TexturesVertex",
             "PerspectiveCameras", "HardPhongShader", "BlendParams"]
    _stub("pytorch3d.renderer", **{n: None for n in names})
    _stub("pytorch3d.renderer.mesh", Textures=None)
    _stub("pytorch3d.structures", Meshes=None)
    _stub("triplane_sampler_cuda")
    _stub("easydict", EasyDict=_EasyDict)
    _stub("tqdm", tqdm=lambda x, *a, **k: x) if "tqdm" not in sys.modules else None

    def fused_leaky_relu(inp, bias=None, negative_slope=0.2, scale=2 ** 0.5):
        if bias is not None:
            inp = inp + bias.view(1, -1, *([1] * (inp.ndim - 2)))
        return torch.nn.functional.leaky_relu(inp, negative_slope) * scale

    class FusedLeakyReLU(torch.nn.Module):
        def __init__(self, channel, bias=True, negative_slope=0.2, scale=2 ** 0.5):
            super().__init__()
            self.bias = torch.nn.Parameter(torch.zeros(channel)) if bias else None
            self.negative_slope, self.scale = negative_slope, scale

        def forward(self, x):
            return fused_leaky_relu(x, self.bias, self.negative_slope, self.scale)

    def upfirdn2d(*a, **k):
        raise NotImplementedError("upfirdn2d is off the hot path")

    _stub("libraries.stylegan2_pytorch.op", FusedLeakyReLU=FusedLeakyReLU, fused_leaky_relu=fused_leaky_relu,
          upfirdn2d=upfirdn2d, conv2d_gradfix=None)
    _stub("op", FusedLeakyReLU=FusedLeakyReLU, fused_leaky_relu=fused_leaky_relu, upfirdn2d=upfirdn2d,
          conv2d_gradfix=None)
    _installed = True


def nerf_config(Nc=48, Nf=48, origin_location="center+head", coordinate_scale=3, constant_triplane=True,
                render_bs=16384):
    return _EasyDict(dict(
        final_activation="tanh", hidden_size=256, out_dim=256, Nc=Nc, Nf=Nf, no_ray_direction=False,
        mask_input=False, origin_location=origin_location, coordinate_scale=coordinate_scale,
        multiply_density_with_triplane_wieght=False, clamp_mask=False, constant_triplane=constant_triplane,
        constant_trimask=False, constant_trimask_lr_mul=1, deformation_field=False, selector_mlp=False,
        no_selector=False, render_bs=render_bs))


def gen_config(**kw):
    return _EasyDict(dict(z_dim=256, patch_size=24, background_ratio=0.7, use_triplane=True,
                          crop_background=False, pretrained_background=False, nerf_params=nerf_config(**kw)))


def build_generator(size, canonical, parents, seed=0, **kw):
    """Random-init reference ``TriNARFGenerator`` (black background, P=24)."""
    install()
    from models.generator import TriNARFGenerator
    torch.manual_seed(seed)
    gen = TriNARFGenerator(gen_config(**kw), size, num_bone=24, num_bone_param=23, parent_id=parents,
                           black_background=True)
    gen.register_canonical_pose(canonical)
    return gen
