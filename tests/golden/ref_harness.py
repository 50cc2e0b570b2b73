"""Load the UNMODIFIED reference so it can pin the oracle and stand beside the kernels as the second oracle / baseline.

The reference is imported from ``/root/reference`` where that exists (the authoring container) and otherwise from
``oracle/_ref/reference_py.zip``, the byte-identical archive of its hot-path Python that ``oracle/build_ref.py`` packs
(git-ignored build output; it travels to the GPU box).  Nothing is copied or edited on disk: a handful of hard-coded
``"cuda"`` device strings are substituted in memory at import time so the same code runs on
``DEVICE`` (CPU here, CUDA on the box), and the missing third-party modules are stubbed.  TEST / BASELINE
INFRASTRUCTURE: imported only by ``tests/`` and ``bench.py``'s reference arm.
"""
import importlib.abc
import importlib.util
import os
import sys
import types
import zipfile

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_ZIP = os.path.join(os.path.dirname(os.path.dirname(_HERE)), "oracle", "_ref", "reference_py.zip")
# ENARF_REF_FROM_ARCHIVE=1 forces the archive even where the live tree exists (to test what the GPU box will see)
REF_ROOT = "/root/reference" if (os.path.isdir("/root/reference/libraries")
                                 and not os.environ.get("ENARF_REF_FROM_ARCHIVE")) else _ZIP
DEVICE = torch.device("cpu")     # where the reference's hard-coded "cuda" tensors go; set_device() changes it

_SUBST = {
    "libraries/NeRF/rendering.py": [
        ('device="cuda"', 'device=_DEV()'),
        ("torch.cuda.FloatTensor(batchsize, 1, n, Nf).uniform_()",
         "_s.modules['ref_harness'].draw_jitter(batchsize, 1, n, Nf)"),
    ],
    "libraries/NeRF/ray_sampler.py": [('device="cuda"', 'device=_DEV()')],
    # the scatter buffer takes the dtype of its operands (float32 in the default run: no change) so that the
    # fp64-cast reference runs
    "libraries/triplane/sampling.py": [("dtype=torch.float32)", "dtype=position.dtype)")],
    # device AND dtype of the pixel grid: identical to ``.float().cuda(...)`` in the default fp32 run, and lets the
    # fp64-cast reference (ref_runner.run(dtype=float64)) keep its intrinsics in double
    "models/generator.py": [("inv_intrinsics.float().cuda(homo_img.device)", "inv_intrinsics.to(homo_img)")],
}
_PRELUDE = "import sys as _s; _DEV = lambda: _s.modules['ref_harness'].DEVICE; "


FORCED_JITTER = None   # tests can pin the fine pass's jitter (ref_runner.run(force_draws=...))


def draw_jitter(*shape):
    """rendering.py:194 (``torch.cuda.FloatTensor(batchsize, 1, n, Nf).uniform_()``) on DEVICE, or the pinned draw."""
    if FORCED_JITTER is not None:
        return FORCED_JITTER.reshape(shape).to(DEVICE)
    return torch.empty(*shape, device=DEVICE).uniform_()


def set_device(device):
    global DEVICE
    DEVICE = torch.device(device)


def _read(rel):
    if os.path.isdir(REF_ROOT):
        with open(os.path.join(REF_ROOT, rel)) as f:
            return f.read()
    with zipfile.ZipFile(REF_ROOT) as z:
        return z.read(rel).decode()


def _exists(rel):
    if os.path.isdir(REF_ROOT):
        return os.path.exists(os.path.join(REF_ROOT, rel))
    with zipfile.ZipFile(REF_ROOT) as z:
        return rel in z.namelist()


class _Loader(importlib.abc.Loader):
    def __init__(self, rel):
        self.rel = rel

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        src = _read(self.rel)
        for a, b in _SUBST[self.rel]:
            assert a in src, (self.rel, a)
            src = src.replace(a, b)
        src = _PRELUDE + src  # same line as the file's first statement: line numbers stay stable
        module.__file__ = os.path.join(REF_ROOT, self.rel)
        exec(compile(src, module.__file__, "exec"), module.__dict__)


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        rel = fullname.replace(".", "/") + ".py"
        if rel in _SUBST and _exists(rel):
            return importlib.util.spec_from_loader(fullname, _Loader(rel))
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
    return os.path.isdir(os.path.join(REF_ROOT, "libraries")) or os.path.isfile(REF_ROOT)


def install():
    """Make ``import models.generator`` etc. resolve to the reference.  Idempotent."""
    global _installed
    if _installed:
        return
    assert available(), "reference not present (neither /root/reference nor oracle/_ref/reference_py.zip)"
    sys.modules.setdefault("ref_harness", sys.modules[__name__])
    # the reference appends "libraries/stylegan2_ada_pytorch" / "libraries/stylegan2_pytorch" relative to its own
    # root to sys.path (triplane_nerf.py:13, custom_stylegan2/net.py:12): give the absolute equivalents instead of
    # changing the working directory (both a directory tree and a zip archive work as import roots)
    sys.path.insert(0, REF_ROOT)
    sys.path.append(os.path.join(REF_ROOT, "libraries", "stylegan2_ada_pytorch"))
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
    ext = None
    if torch.cuda.is_available():     # the reference's own extension, compiled unmodified by oracle/build_ref.py
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(_HERE)), "oracle"))
        import build_ref
        ext = build_ref.load()
    if ext is not None:
        sys.modules["triplane_sampler_cuda"] = ext
    else:
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

    def upfirdn2d(x, kernel, up=1, down=1, pad=(0, 0)):
        """Pure-torch stand-in for the StyleGAN2 op the discriminator's Blur layers call (the original JIT-compiles
        CUDA at import): zero-stuff by ``up``, pad, correlate every channel with the flipped 2-D FIR, keep every
        ``down``-th sample."""
        B, C, H, W = x.shape
        if up > 1:
            z = x.new_zeros(B, C, H, up, W, up)
            z[:, :, :, 0, :, 0] = x
            x = z.reshape(B, C, H * up, W * up)
        x = torch.nn.functional.pad(x, (pad[0], pad[1], pad[0], pad[1]))
        k = torch.flip(kernel, (0, 1)).to(x)[None, None].expand(C, 1, -1, -1)
        x = torch.nn.functional.conv2d(x, k, groups=C)
        return x[:, :, ::down, ::down]

    gradfix = types.SimpleNamespace(conv2d=torch.nn.functional.conv2d,
                                    conv_transpose2d=torch.nn.functional.conv_transpose2d)

    _stub("libraries.stylegan2_pytorch.op", FusedLeakyReLU=FusedLeakyReLU, fused_leaky_relu=fused_leaky_relu,
          upfirdn2d=upfirdn2d, conv2d_gradfix=gradfix)
    _stub("op", FusedLeakyReLU=FusedLeakyReLU, fused_leaky_relu=fused_leaky_relu, upfirdn2d=upfirdn2d,
          conv2d_gradfix=gradfix)
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
