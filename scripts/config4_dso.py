"""BASELINE configs[3] (DSO / denarf step: B=4, 512x512 rays per sample, 96+96 samples per ray, bf16 tri-planes)
through the public mirror: DSONARFGenerator.forward (mask-based ray selection, time/pose latents, deformation-field
producer, sparse render) + PhotometricLoss + backward, CUDA-event timed per stage.  Two producers are compared:

  handoff : enarf_warp_planes_fwd/bwd writing / reading the renderer's channels-last planes
  nchw    : the reference's formulation (F.grid_sample + cat in eager torch, models/narf.py:43-57) handed to the
            renderer as an NCHW tri-plane, which then re-packs it and un-packs the gradient

    python scripts/config4_dso.py            # env: B, SIZE, NS (samples per pass), STEPS
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from enarf_gan_b200 import synthetic as S
from enarf_gan_b200.generator import DSONARFGenerator, PhotometricLoss

B, size = int(os.environ.get("B", 4)), int(os.environ.get("SIZE", 512))
NS, steps = int(os.environ.get("NS", 96)), int(os.environ.get("STEPS", 5))
dev = torch.device("cuda:0")


class Flow(torch.nn.Module):
    """Stand-in for the StyleGAN2-ADA flow head (out of scope): a learnable flow field."""

    def __init__(self, f):
        super().__init__()
        self.f = torch.nn.Parameter(f)

    def forward(self, z, bone_length, truncation_psi=1):
        return self.f[:z.shape[0]]


class TorchWarp(torch.nn.Module):
    """The reference's producer in eager torch (narf.py:43-57)."""

    def __init__(self, canon, f):
        super().__init__()
        self.tri_plane = torch.nn.Parameter(canon)
        self.f = torch.nn.Parameter(f)

    def forward(self, z, bone_length, truncation_psi=1):
        bs, S_ = z.shape[0], 256
        fl = self.f[:bs].reshape(bs * 3, 2, S_, S_).permute(0, 2, 3, 1)
        ar = torch.arange(S_, device=z.device)
        grid = torch.stack(torch.meshgrid(ar, ar, indexing="ij")[::-1], dim=2) + 0.5
        grid = (grid + fl) / 128 - 1
        tp = self.tri_plane.expand(bs, -1, -1, -1)
        w = F.grid_sample(tp[:, :96].reshape(bs * 3, 32, S_, S_), grid, align_corners=False)
        return torch.cat([w.reshape(bs, 96, S_, S_), tp[:, 96:]], dim=1)


def build(mode, canon, flow):
    nerf_params = dict(Nc=NS, Nf=NS, origin_location="center_fixed", coordinate_scale=3, no_ray_direction=True,
                       time_conditional=True, pose_conditional=True, render_bs=16384,
                       deformation_field=(mode == "handoff"))
    config = dict(use_triplane=True, ray_batchsize=size * size, nerf_params=nerf_params)
    torch.manual_seed(1238)          # same random-init decoder for both producers
    kw = dict(flow_generator=Flow(flow)) if mode == "handoff" else dict(tri_plane_gen=TorchWarp(canon, flow))
    gen = DSONARFGenerator(config, size, num_bone=24, parent_id=S.SMPL_PARENTS, num_bone_param=23,
                           plane_dtype=torch.bfloat16, **kw)
    gen.register_canonical_pose(S.canonical_pose())
    if mode == "handoff":
        with torch.no_grad():
            gen.nerf.tri_plane.copy_(canon)
    gen.nerf.rng_mode = "philox"
    return gen.to(dev).train()


def main():
    inp = {k: v.to(dev) for k, v in S.make_inputs(B, size, seed=1238).items()}
    canon = S.make_triplane(1, num_part=23, seed=1238)
    flow = S.make_flow(B, seed=1238)
    fg = S.make_foreground_mask(B, size, seed=1238).to(dev)
    t = torch.rand(B, device=dev)
    gt_color = torch.rand(B, 3, size, size, device=dev) * 2 - 1
    loss_fn = PhotometricLoss(dict(nerf_loss_type="mae", color_coef=1, mask_coef=1))
    for mode in ("handoff", "nchw"):
        gen = build(mode, canon, flow)
        torch.cuda.reset_peak_memory_stats()
        rows = []
        for it in range(steps):
            for p in gen.parameters():
                p.grad = None
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            torch.cuda.synchronize()
            e[0].record()
            color, mask, grid = gen(inp["pose_to_camera"], None, fg, t, inp["bone_length"], inp["inv_intrinsics"])
            lc, lm = loss_fn(grid, color, mask, gt_color, fg)
            e[1].record()
            (lc + lm).backward()
            e[2].record()
            torch.cuda.synchronize()
            rows.append((e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])))
        fwd, bwd = sorted(r[0] for r in rows[1:])[len(rows[1:]) // 2], sorted(r[1] for r in rows[1:])[len(rows[1:]) // 2]
        R = B * size * size
        tp_grad = gen.nerf.tri_plane.grad if mode == "handoff" else gen.nerf.tri_plane_gen.tri_plane.grad
        f_grad = gen.nerf.flow_generator.f.grad if mode == "handoff" else gen.nerf.tri_plane_gen.f.grad
        print(json.dumps({"producer": mode, "B": B, "rays": R, "samples_per_ray": 2 * NS, "planes": "bf16",
                          "fwd_ms": round(fwd, 2), "bwd_ms": round(bwd, 2), "step_ms": round(fwd + bwd, 2),
                          "rays_per_s": round(R / (fwd + bwd) * 1e3), "points_per_s": round(R * 2 * NS / (fwd + bwd) * 1e3),
                          "loss": round(float((lc + lm).detach()), 5), "mask_mean": round(float(mask.mean()), 4),
                          "d_triplane_absmax": float(tp_grad.abs().max()), "d_flow_absmax": float(f_grad.abs().max()),
                          "peak_mem_GB": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}))
        del gen
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
