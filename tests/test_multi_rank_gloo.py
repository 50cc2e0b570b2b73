"""World-size-2 gloo test of the N>1 host logic (runs on CPU): sample sharding and the batch-global near/far
exchange that keeps a ray-sharded render identical to the single-process one."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, pose, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from enarf_gan_b200 import rendering
        lo, hi = rendering.shard_samples(pose.shape[0], rank, world)
        nf = rendering.global_near_far(pose[lo:hi])
        torch.save({"range": (lo, hi), "nearfar": nf}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_shard_samples_partitions_the_batch():
    from enarf_gan_b200 import rendering
    for B in (1, 2, 7, 8, 32):
        for world in (1, 2, 3, 8):
            blocks = [rendering.shard_samples(B, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
            sizes = [b[1] - b[0] for b in blocks]
            assert max(sizes) - min(sizes) <= 1


def test_global_near_far_matches_single_process(tmp_path):
    import enarf_oracle as O
    from enarf_gan_b200 import synthetic as S
    inp = S.make_inputs(5, 16, seed=77)
    pose, _ = O.transform_pose(inp["pose_to_camera"], inp["bone_length"], "center+head", S.SMPL_PARENTS)
    pose = pose.clone()
    pose[:, :, :3, 3] *= 3          # coordinate_scale (rendering.py:258-260)
    near, far = O.near_far(pose)
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), pose, str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(world)]
    assert got[0]["range"] == (0, 3) and got[1]["range"] == (3, 5)
    for g in got:   # every rank holds the full-batch scalars, bit-identical to the oracle's
        assert torch.equal(g["nearfar"], torch.stack([near, far]).float())


def _gather_worker(rank, world, port, n, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from enarf_gan_b200 import rendering
        lo, hi = rendering.shard_rays(n, rank, world)
        ray = torch.arange(lo, hi, dtype=torch.float32)
        local = torch.stack([ray, ray * 2, ray * 3])                     # a (3, n_local) "colour" block
        full = rendering.gather_rays(local, n)
        mask = rendering.gather_rays(ray + 0.5, n)
        torch.save({"full": full, "mask": mask}, os.path.join(out_dir, f"g{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_ray_sharding_gathers_the_whole_image(tmp_path):
    """B < G: one image's rays split over two ranks (uneven block sizes) and all-gathered back in ray order."""
    n, world = 17 * 17, 2            # odd ray count: the blocks differ by one ray and are padded for the all_gather
    mp.spawn(_gather_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    ray = torch.arange(n, dtype=torch.float32)
    for r in range(world):
        got = torch.load(os.path.join(tmp_path, f"g{r}.pt"))
        assert torch.equal(got["full"], torch.stack([ray, ray * 2, ray * 3]))
        assert torch.equal(got["mask"], ray + 0.5)


# ---- B < G with gradients: the autograd collectives of rendering.render_entire_img_sharded ------------------------
def _toy_render(planes, weight, lo, hi):
    """A differentiable stand-in for a ray block's render: rays lo..hi read the shared planes / weights."""
    rays = torch.arange(lo, hi, dtype=torch.float32)
    return torch.sin(rays[None] * planes.sum() * 0.01) * weight.sum() + planes[0, lo:hi] * weight[0]


def _grad_worker(rank, world, port, n, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from enarf_gan_b200 import rendering
        torch.manual_seed(3)
        planes = torch.randn(1, n).requires_grad_()                # the owner's values count; others pass same-shaped junk
        weight = torch.randn(2).requires_grad_()
        if rank != 0:
            planes = torch.zeros(1, n).requires_grad_()
            weight = torch.zeros(2).requires_grad_()
        p, w = rendering.from_owner(planes, 0), rendering.from_owner(weight, 0)
        lo, hi = rendering.shard_rays(n, rank, world)
        img = rendering._GatherRays.apply(_toy_render(p, w, lo, hi), n, None)
        (img * torch.arange(n, dtype=torch.float32)).sum().backward()      # the same loss on every rank
        torch.save({"img": img.detach(), "d_planes": planes.grad, "d_weight": weight.grad}, os.path.join(out_dir, f"g{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_ray_sharded_gradients_sum_onto_the_owner(tmp_path):
    n = 37                                           # uneven blocks
    mp.spawn(_grad_worker, args=(2, _free_port(), n, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"g{r}.pt")) for r in range(2))
    torch.manual_seed(3)
    planes = torch.randn(1, n).requires_grad_()
    weight = torch.randn(2).requires_grad_()
    img = _toy_render(planes, weight, 0, n)
    (img * torch.arange(n, dtype=torch.float32)).sum().backward()
    assert torch.allclose(r0["img"], img.detach(), atol=1e-6) and torch.allclose(r1["img"], img.detach(), atol=1e-6)
    assert torch.allclose(r0["d_planes"], planes.grad, rtol=1e-5, atol=1e-5), "owner: sum of both blocks' plane gradients"
    assert torch.allclose(r0["d_weight"], weight.grad, rtol=1e-5, atol=1e-4)
    assert r1["d_planes"].abs().sum() == 0 and r1["d_weight"].abs().sum() == 0, "non-owners receive no gradient"
