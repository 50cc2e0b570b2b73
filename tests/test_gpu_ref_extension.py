"""Row a9 against the reference's OWN CUDA kernels: ``oracle/_ref/triplane_sampler_cuda.so`` is the reference's
cuda_extension compiled unmodified (``oracle/build_ref.py``); on the GPU its forward / backward run next to
``enarf_triplane_sample_fwd/bwd`` on the same inputs.  Skipped when the module was not built (no reference sources)."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import build_ref  # noqa: E402

pytestmark = pytest.mark.gpu
MODES = {"bilinear": 0, "nearest": 1}
PADS = {"zeros": 0, "border": 1, "reflection": 2}


@pytest.fixture(scope="module")
def ref_ext():
    if build_ref.so_path() is None:
        pytest.skip("oracle/_ref/triplane_sampler_cuda.so not built (reference sources absent at build time)")
    return build_ref.load()


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("C", [8, 32])          # 32 = the channels-last fast path, 8 = the generic NCHW kernels
@pytest.mark.parametrize("padding", ["zeros", "border", "reflection"])
@pytest.mark.parametrize("align", [False, True])
def test_bilinear_forward_and_backward_match_reference_kernels(ref_ext, C, padding, align):
    from enarf_gan_b200.triplane_sampler import triplane_sampler
    torch.manual_seed(3)
    N, H, W, n = 2, 33, 47, 4001
    inp = torch.randn(N, 3 * C, H, W, device="cuda", requires_grad=True)
    grid = (torch.rand(N, n, 1, 3, device="cuda") * 2.6 - 1.3).requires_grad_()
    out = triplane_sampler(inp, grid, "bilinear", padding, align)
    ref = ref_ext.triplane_sampler_forward(inp.detach(), grid.detach(), 0, PADS[padding], align)
    assert out.shape == ref.shape
    assert rel(out, ref) < 1e-6
    w = torch.randn_like(ref)
    gi, gg = torch.autograd.grad((out * w).sum(), [inp, grid])
    ri, rg = ref_ext.triplane_sampler_backward(w, inp.detach(), grid.detach(), 0, PADS[padding], align, [True, True])
    assert rel(gi, ri) < 1e-5          # float atomics: summation order differs
    assert rel(gg, rg) < 1e-5


def test_nearest_mode_divergence_is_the_documented_one(ref_ext):
    """The reference's nearest branch ASSIGNS per plane (TriplaneSampler_kernel.cu:79-88), so its output is the last
    plane's sample alone and it rounds half away from zero; ours sums the three planes like the bilinear branch and
    like three F.grid_sample calls (tests/test_gpu_api.py).  No caller of the reference uses nearest."""
    import torch.nn.functional as F
    torch.manual_seed(5)
    N, C, H, W, n = 1, 4, 16, 16, 2000
    inp = torch.randn(N, 3 * C, H, W, device="cuda")
    grid = torch.rand(N, n, 1, 3, device="cuda") * 1.8 - 0.9
    ref = ref_ext.triplane_sampler_forward(inp, grid, 1, 0, False)
    last = F.grid_sample(inp[:, 2 * C:], torch.stack([grid[..., 2], grid[..., 0]], dim=-1), mode="nearest",
                         padding_mode="zeros", align_corners=False)
    assert (ref - last).abs().max().item() < 1e-6 or ((ref - last).abs() > 1e-6).float().mean().item() < 1e-3


def test_reference_shape_of_the_generator_call(ref_ext):
    """The shape the reference's sample_feature uses (sampling.py:22-26): input (1,96,256,256), grid (1,n,1,3)."""
    from enarf_gan_b200.triplane_sampler import triplane_sampler
    torch.manual_seed(4)
    inp = torch.randn(1, 96, 256, 256, device="cuda")
    grid = torch.rand(1, 200000, 1, 3, device="cuda") * 2.2 - 1.1
    out = triplane_sampler(inp, grid)
    ref = ref_ext.triplane_sampler_forward(inp, grid, 0, 0, False)
    assert rel(out, ref) < 1e-6
