"""Time the deformation-field producer (enarf_warp_planes_fwd/bwd) at the DSO batch sizes and, beside it, the
reference's formulation of the same step in eager torch on the same GPU (models/narf.py:43-57: F.grid_sample of the
NCHW planes + cat; its autograd backward), plus the repack / unpack the hand-off layout makes unnecessary.
CUDA events on torch's current stream, 5 warm-up + 20 timed launches, inputs (B x 25 MB out) larger than... no: the
126 MB L2 is flushed between launches by writing a 256 MB buffer.

    python scripts/warp_time.py [B ...]
"""
import json
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from enarf_gan_b200 import ops, synthetic as S  # noqa: E402


def timed(fn, flush, n=20, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def reference_warp(tri_plane, flow):
    bs, S_ = flow.shape[0], 256
    fl = flow.reshape(bs * 3, 2, S_, S_).permute(0, 2, 3, 1)
    ar = torch.arange(S_, device=flow.device)
    grid = torch.stack(torch.meshgrid(ar, ar, indexing="ij")[::-1], dim=2) + 0.5
    grid = (grid + fl) / 128 - 1
    tp = tri_plane.expand(bs, -1, -1, -1)
    w = F.grid_sample(tp[:, :96].reshape(bs * 3, 32, S_, S_), grid, align_corners=False)
    return torch.cat([w.reshape(bs, 96, S_, S_), tp[:, 96:]], dim=1)


def main():
    peaks = json.load(open("MEASURED_PEAKS.json")) if len(sys.argv) >= 0 else {}
    peak = float(peaks.get("hbm_gbs", 0) or 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    canon = S.make_triplane(1, num_part=23, seed=1).cuda()
    cl = ops.repack_planes(canon)[0]
    for B in [int(a) for a in sys.argv[1:]] or [4, 16]:
        flow = S.make_flow(B, seed=2).cuda()
        texels = B * 3 * 256 * 256
        row = {"B": B}
        for dt, name, s in ((torch.float32, "f32", 4), (torch.bfloat16, "bf16", 2)):
            out = ops.warp_planes(cl, flow, dt)
            ms = timed(lambda: ops.warp_planes(cl, flow, dt), flush)
            alg = texels * (8 + 32 * s) + 3 * 256 * 256 * 128            # flow + output line, canonical planes once
            row[f"fwd_{name}_ms"] = round(ms, 4)
            row[f"fwd_{name}_GBps"] = round(alg / ms / 1e6, 1)
            g = torch.randn_like(out)
            d_c, d_f = torch.zeros_like(cl), torch.empty_like(flow)

            def bwd():
                d_c.zero_()
                ops.check(ops.lib().enarf_warp_planes_bwd(g.data_ptr(), ops._dtype_code(dt), cl.data_ptr(), flow.data_ptr(),
                                                          B, 256, 256, d_c.data_ptr(), d_f.data_ptr(), ops.stream_ptr()))
            ms = timed(bwd, flush)
            alg = texels * (8 + 32 * s + 8) + 2 * 3 * 256 * 256 * 128    # flow + gradient line + d_flow, canon + d_canon
            row[f"bwd_{name}_ms"] = round(ms, 4)
            row[f"bwd_{name}_GBps"] = round(alg / ms / 1e6, 1)
        # the reference's formulation in eager torch, and the layout conversions the hand-off removes
        tp = canon.clone().requires_grad_()
        fl = flow.clone().requires_grad_()
        row["torch_ref_fwd_ms"] = round(timed(lambda: reference_warp(tp.detach(), flow), flush), 4)
        up = torch.randn(B, 165, 256, 256, device="cuda")

        def ref_fb():
            tp.grad = fl.grad = None
            reference_warp(tp, fl).backward(up)
        row["torch_ref_fwd_bwd_ms"] = round(timed(ref_fb, flush), 4)
        full = reference_warp(canon, flow)
        row["repack_ms"] = round(timed(lambda: ops.repack_planes(full), flush), 4)
        if peak:
            row["hbm_peak_GBps"] = peak
        print(json.dumps(row))


if __name__ == "__main__":
    main()
