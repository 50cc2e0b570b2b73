"""Raw-op microbench of SURVEY 8a row a9 / 8d config 2: the stand-alone 3-plane sampler on the shape the reference's
generator would feed it (input (1,96,256,256), grid (1,N,1,3), N = 128*128*8*48 = 6 291 456 points), ours
(enarf_triplane_sample_fwd/bwd through the reference-facing ``triplane_sampler`` Function) next to the reference's own
CUDA extension compiled unmodified (oracle/_ref, built by oracle/build_ref.py).  CUDA events, 10 warm-up + 50 timed.
The reference launches on the legacy default stream (TriplaneSampler_kernel.cu:246,286), which is torch's current
stream here, so one pair of CUDA events brackets both implementations the same way.

    python scripts/raw_op_bench.py [N]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch

import build_ref
from enarf_gan_b200._abi import check, lib, stream_ptr


def timed(fn, warm=10, n=50):
    """Mean ms per call: CUDA events on torch's current stream, which here is the legacy default stream -- the one the
    reference's kernels are launched on -- with a device synchronize on both sides."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 128 * 128 * 8 * 48
    C, H, W = 32, 256, 256
    torch.manual_seed(0)
    inp = torch.randn(1, 3 * C, H, W, device="cuda")
    grid = torch.rand(1, N, 1, 3, device="cuda") * 2.2 - 1.1
    gout = torch.randn(1, C, N, 1, device="cuda")
    out = torch.empty(1, C, N, 1, device="cuda")
    g_in, g_grid = torch.zeros_like(inp), torch.zeros_like(grid)
    L = lib()

    from enarf_gan_b200 import ops
    feat = ops.repack_planes(inp)
    d_feat = torch.zeros_like(feat)

    def ours_fwd():          # what triplane_sampler() does for C == 32: repack to channels-last, then gather
        f = ops.repack_planes(inp)
        check(L.enarf_triplane_sample_cl_fwd(f.data_ptr(), grid.data_ptr(), 1, H, W, N, 0, 0, out.data_ptr(), stream_ptr()))

    def ours_bwd():          # zero-fill, scatter + grid gradient, transpose back to the NCHW grad_input
        d_feat.zero_()
        g_grid.zero_()
        check(L.enarf_triplane_sample_cl_bwd(gout.data_ptr(), feat.data_ptr(), grid.data_ptr(), 1, H, W, N, 0, 0,
                                             d_feat.data_ptr(), g_grid.data_ptr(), stream_ptr()))
        ops.unpack_plane_grads(d_feat, g_in, accumulate=False)

    def generic_fwd():       # the NCHW kernel kept for other channel counts / nearest
        check(L.enarf_triplane_sample_fwd(inp.data_ptr(), grid.data_ptr(), 1, C, H, W, N, 1, 0, 0, 0, out.data_ptr(),
                                          stream_ptr()))

    row = {"op": "triplane_sampler", "input": [1, 3 * C, H, W], "points": N,
           "ours_fwd_ms": round(timed(ours_fwd), 4), "ours_bwd_ms": round(timed(ours_bwd), 4),
           "ours_generic_nchw_fwd_ms": round(timed(generic_fwd), 4)}
    ours_fwd()
    # algorithmic bytes: 12 B grid + 12 texels x 32 channels x 4 B gathered + 128 B output per point
    row["ours_fwd_gather_GBps"] = round(N * (12 + 12 * 128 + 128) / row["ours_fwd_ms"] / 1e6, 1)
    ref = build_ref.load()
    if ref is not None:
        ro = ref.triplane_sampler_forward(inp, grid, 0, 0, False)
        row["max_abs_diff_fwd"] = float((ro - out).abs().max())
        row["ref_fwd_ms"] = round(timed(lambda: ref.triplane_sampler_forward(inp, grid, 0, 0, False)), 4)
        row["ref_bwd_ms"] = round(timed(lambda: ref.triplane_sampler_backward(gout, inp, grid, 0, 0, False, [True, True])), 4)
        row["speedup_fwd"] = round(row["ref_fwd_ms"] / row["ours_fwd_ms"], 2)
        row["speedup_bwd"] = round(row["ref_bwd_ms"] / row["ours_bwd_ms"], 2)
    else:
        row["ref"] = "oracle/_ref not built"
    print(json.dumps(row))


if __name__ == "__main__":
    main()
