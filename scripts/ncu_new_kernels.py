"""Launch the round's new small kernels a few times each (ncu target): warp producer fwd/bwd at B=4 (fp32, bf16 out),
channels-last sampler fwd/bwd at the config-2 raw-op shape."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from enarf_gan_b200 import ops, synthetic as S
from enarf_gan_b200._abi import check, lib, stream_ptr

canon = S.make_triplane(1, num_part=23, seed=1).cuda()
cl = ops.repack_planes(canon)[0]
flow = S.make_flow(4, seed=2).cuda()
N = 128 * 128 * 8 * 48
inp = torch.randn(1, 96, 256, 256, device="cuda")
feat = ops.repack_planes(inp)
grid = torch.rand(1, N, 1, 3, device="cuda") * 2.2 - 1.1
gout = torch.randn(1, 32, N, 1, device="cuda")
out = torch.empty(1, 32, N, 1, device="cuda")
d_feat, g_grid = torch.zeros_like(feat), torch.zeros_like(grid)
L = lib()
for it in range(3):
    for dt in (torch.float32, torch.bfloat16):
        o = ops.warp_planes(cl, flow, dt)
        d_c, d_f = torch.zeros_like(cl), torch.empty_like(flow)
        g = torch.randn_like(o)
        check(L.enarf_warp_planes_bwd(g.data_ptr(), ops._dtype_code(dt), cl.data_ptr(), flow.data_ptr(), 4, 256, 256,
                                      d_c.data_ptr(), d_f.data_ptr(), stream_ptr()))
    check(L.enarf_triplane_sample_cl_fwd(feat.data_ptr(), grid.data_ptr(), 1, 256, 256, N, 0, 0, out.data_ptr(), stream_ptr()))
    check(L.enarf_triplane_sample_cl_bwd(gout.data_ptr(), feat.data_ptr(), grid.data_ptr(), 1, 256, 256, N, 0, 0,
                                         d_feat.data_ptr(), g_grid.data_ptr(), stream_ptr()))
torch.cuda.synchronize()
print("ok")
