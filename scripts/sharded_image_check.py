"""B < G on real GPUs: one image's rays split over the ranks of a torchrun job (NCCL), every rank renders its
block with the replicated model and the blocks are all-gathered; rank 0 compares the assembled image with its own
one-shot render of the same scene and reports the timings.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
        scripts/sharded_image_check.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from enarf_gan_b200 import rendering, synthetic as S
from enarf_gan_b200.generator import TriNARFGenerator

SIZE = int(os.environ.get("SIZE", 256))


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dev = torch.device("cuda")
    if world > 1:
        dist.init_process_group("nccl")
    torch.manual_seed(0)                                      # replicated random-init decoder on every rank
    cfg = dict(z_dim=256, background_ratio=0.7,
               nerf_params=dict(Nc=48, Nf=48, origin_location="center+head", coordinate_scale=3, render_bs=1 << 20,
                                constant_triplane=True))
    gen = TriNARFGenerator(cfg, SIZE, num_bone=24, num_bone_param=23, parent_id=S.SMPL_PARENTS, black_background=True)
    gen.register_canonical_pose(S.canonical_pose())
    with torch.no_grad():
        gen.nerf.tri_plane.copy_(S.make_triplane(1, seed=5))
    gen = gen.to(dev).eval()
    nerf = gen.nerf
    nerf.rng_mode = "philox"
    inp = {k: v.to(dev) for k, v in S.make_inputs(1, SIZE, seed=5).items()}
    z = S.make_latent(1, seed=5).to(dev)
    pose, bl = nerf.transform_pose(inp["pose_to_camera"], inp["bone_length"])
    mi = {"z": z[:, :512], "z_rend": z[:, 512:], "bone_length": bl, "truncation_psi": 1,
          "tri_plane_feature": nerf.compute_tri_plane_feature(z[:, :512], bl)}

    def timed(fn, n=5):
        fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            out = fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / n], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return out, t.item()

    sharded, ms_sh = timed(lambda: rendering.render_entire_img_sharded(nerf, pose, inp["inv_intrinsics"], render_size=SIZE,
                                                                      Nc=48, Nf=48, model_input=mi))
    full, ms_full = timed(lambda: rendering.render_entire_img(nerf, pose, inp["inv_intrinsics"], render_size=SIZE, Nc=48,
                                                              Nf=48, model_input=mi))
    again, _ = timed(lambda: rendering.render_entire_img(nerf, pose, inp["inv_intrinsics"], render_size=SIZE, Nc=48, Nf=48,
                                                         model_input=mi), n=1)
    if rank == 0:
        (c1, m1, d1), (c0, m0, d0) = sharded, full
        noise = float((again[1] - m0).abs().mean())         # two one-shot renders with independent fine-pass draws
        print(json.dumps({"world": world, "image": [SIZE, SIZE], "ms_sharded": round(ms_sh, 3), "ms_one_gpu": round(ms_full, 3),
                          "mask_mean_abs_diff": float((m1 - m0).abs().mean()), "mask_mean_abs_diff_between_two_one_gpu_renders": noise, "color_mean_abs_diff": float((c1 - c0).abs().mean()),
                          "mask_mean": float(m0.mean()), "valid_fraction": float((m0 > 0).float().mean())}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
