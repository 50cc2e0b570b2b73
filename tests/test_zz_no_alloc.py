"""enarf_ray_scores_static (csrc/dso.cu): the allocation-free variant of enarf_ray_scores -- its row-pass scratch is a
file-scope `static __device__` array inside the library instead of a caller buffer.  It has no counterpart in the
reference; the checker is the existing entry point, called through the
same C ABI on the same inputs.  The two run identical kernels, so the bar is bit-exactness (the DSO tests' own bar for
the score image's consumer, the selected ray set).  Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

LIMIT = 1 << 20     # ENARF_STATIC_SCRATCH_FLOATS in include/enarf_b200.h


def _mask(B, h, w, seed):
    g = torch.Generator().manual_seed(seed)
    m = (torch.rand(B, h, w, generator=g) > 0.995).float()
    return m.cuda()


def _both(lib, abi, mask, radius, noise, seed):
    B, h, w = mask.shape
    st = abi.stream_ptr()
    nz = abi.ptr(noise, allow_none=True)
    scratch = torch.empty_like(mask)
    ref = torch.full_like(mask, float("nan"))
    abi.check(lib.enarf_ray_scores(abi.ptr(mask), B, h, w, radius, nz, seed, abi.ptr(scratch), abi.ptr(ref), st))
    out = torch.full_like(mask, float("nan"))
    abi.check(lib.enarf_ray_scores_static(abi.ptr(mask), B, h, w, radius, nz, seed, abi.ptr(out), st))
    torch.cuda.synchronize()
    return out, ref


def test_static_scratch_variant_matches_the_allocating_entry_point_twice_in_a_row():
    from enarf_gan_b200 import _abi as abi
    lib = abi.lib()
    torch.cuda.set_device(0)
    # two calls in a row with different shapes / contents: the second must not see anything of the first in the
    # shared array (a large image first, then a smaller one that covers only part of what the first call wrote)
    cases = [(4, 512, 512, 64, True, 11), (3, 96, 80, 5, False, 12), (4, 512, 512, 64, False, 13)]
    for B, h, w, radius, given_noise, seed in cases:
        mask = _mask(B, h, w, seed)
        noise = torch.rand(B, h, w, generator=torch.Generator().manual_seed(seed + 100)).cuda() if given_noise else None
        out, ref = _both(lib, abi, mask, radius, noise, seed)
        assert torch.isfinite(ref).all()
        assert torch.equal(out, ref), (B, h, w, (out - ref).abs().max().item())


def test_static_scratch_variant_allocates_nothing_and_rejects_oversize():
    from enarf_gan_b200 import _abi as abi
    lib = abi.lib()
    torch.cuda.set_device(0)
    mask = _mask(2, 128, 128, 3)
    out = torch.empty_like(mask)
    st = abi.stream_ptr()
    abi.check(lib.enarf_ray_scores_static(abi.ptr(mask), 2, 128, 128, 8, None, 5, abi.ptr(out), st))   # set-up call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(8):
        abi.check(lib.enarf_ray_scores_static(abi.ptr(mask), 2, 128, 128, 8, None, 5, abi.ptr(out), st))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "the call must not allocate device memory"
    # beyond the limit: a clear error, nothing launched (the pointers are never dereferenced)
    rc = lib.enarf_ray_scores_static(abi.ptr(mask), 1, 1024, 1025, 8, None, 5, abi.ptr(out), st)
    assert rc == 2      # ENARF_ERR_UNSUPPORTED
    msg = lib.enarf_last_error().decode()
    assert "ENARF_STATIC_SCRATCH_FLOATS" in msg and str(LIMIT) in msg, msg
    # exactly at the limit is accepted
    big = _mask(1, 1024, 1024, 4)
    o, r = _both(lib, abi, big, 3, None, 9)
    assert torch.equal(o, r)
