"""World-size-2 gloo test (CPU) of the training-step plumbing: with the batch sample-sharded over two
ranks and generator / discriminator under DDP, one ``train.train_step`` leaves every rank with the gradients -- and,
after the optimiser step, the parameters -- that a single process gets from averaging the two shards' losses; gradient
accumulation (n_accum_step = 2) crosses the wire once.  The renderer needs a GPU, so a small torch module with the
generator's call signature stands in for it here; the GPU test of the same step with the real kernels is
tests/test_gpu_train.py."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SIZE, ZT = 8, 12


class ToyGenerator(nn.Module):
    """forward(pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics) -> (img, mask, None, None)."""
    background_ratio = 0.3

    def __init__(self):
        super().__init__()
        self.lin = nn.Linear(ZT + 23, 4 * SIZE * SIZE)

    def forward(self, pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics):
        h = self.lin(torch.cat([z, bone_length[:, :, 0]], dim=1)).reshape(-1, 4, SIZE, SIZE)
        return torch.tanh(h[:, :3]), torch.sigmoid(h[:, 3]), None, None


class ToyDiscriminator(nn.Module):
    def __init__(self):
        super().__init__()
        self.net = nn.Sequential(nn.Conv2d(3, 4, 3, padding=1), nn.Softplus(), nn.Flatten(), nn.Linear(4 * SIZE * SIZE, 1))

    def forward(self, img, ddp=False, world_size=1):
        return self.net(img)


def _batch(B):
    g = torch.Generator().manual_seed(11)
    return dict(pose_to_camera=torch.randn(B, 24, 4, 4, generator=g), pose_to_world=torch.randn(B, 24, 4, 4, generator=g),
                bone_length=torch.rand(B, 23, 1, generator=g), inv_intrinsic=torch.randn(B, 3, 3, generator=g),
                bone_mask=(torch.rand(B, SIZE, SIZE, generator=g) > 0.5).float(), real_img=torch.randn(B, 3, SIZE, SIZE, generator=g))


def _models():
    torch.manual_seed(5)
    return ToyGenerator(), ToyDiscriminator()


def _run_shard(gen, dis, data, lo, hi, seed, n_accum, ddp=False, world=1):
    from enarf_gan_b200 import train
    cfg = train.StepConfig(z_total=ZT, n_accum_step=n_accum, adv_loss_type="ce", r1_every=16)
    go = torch.optim.SGD(gen.parameters(), lr=0.1)
    do = torch.optim.SGD(dis.parameters(), lr=0.1)
    d = {k: v[lo:hi] for k, v in data.items()}
    zg = torch.Generator().manual_seed(seed)
    train.train_step(1, hi - lo, gen, dis, go, do, cfg, d["pose_to_camera"], d["pose_to_world"], d["bone_length"],
                     d["inv_intrinsic"], d["bone_mask"], d["real_img"], ddp=ddp, world_size=world, generator=zg)
    return gen, dis


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_accum, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from enarf_gan_b200 import rendering, train
        data = _batch(4 * world)
        gen, dis = _models()
        gen, dis = train.wrap_ddp(gen), train.wrap_ddp(dis)
        assert isinstance(gen, nn.parallel.DistributedDataParallel)
        lo, hi = rendering.shard_samples(4 * world, rank, world)
        _run_shard(gen, dis, data, lo, hi, 100 + rank, n_accum, ddp=True, world=world)
        torch.save({"gen": gen.module.state_dict(), "dis": dis.module.state_dict()}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _single_process_reference(n_accum, world=2):
    """What DDP must reproduce: one SGD step on the MEAN of the per-shard gradients (same latent seeds per shard).  The
    discriminator phase sees the generator after its step, exactly as inside train_step."""
    from enarf_gan_b200 import rendering
    data = _batch(4 * world)
    shards = []
    for r in range(world):
        gen, dis = _models()
        p0 = [p.detach().clone() for p in list(gen.parameters()) + list(dis.parameters())]
        lo, hi = rendering.shard_samples(4 * world, r, world)
        _run_shard(gen, dis, data, lo, hi, 100 + r, n_accum)
        shards.append([p.detach() for p in list(gen.parameters()) + list(dis.parameters())])
    # SGD: p1 = p0 - lr * g, so averaging the stepped parameters averages the gradients -- valid for the generator,
    # whose gradient is computed at p0 on both shards
    return p0, [torch.stack(ps).mean(0) for ps in zip(*shards)]


def _check(tmp_path, n_accum):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_accum, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(2))
    for k in r0["gen"]:
        assert torch.equal(r0["gen"][k], r1["gen"][k]), f"ranks disagree on generator {k}"
    for k in r0["dis"]:
        assert torch.equal(r0["dis"][k], r1["dis"][k]), f"ranks disagree on discriminator {k}"
    p0, mean = _single_process_reference(n_accum)
    gen, _ = _models()
    n_gen = len(list(gen.parameters()))
    got = list(r0["gen"].values())
    for i, (a, b, c) in enumerate(zip(got, mean[:n_gen], p0[:n_gen])):
        assert not torch.equal(a, c), "the step must move the generator"
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6), f"generator parameter {i}: DDP != mean of the shard gradients"


def test_ddp_step_equals_mean_of_shard_gradients(tmp_path):
    _check(tmp_path, n_accum=1)


def test_gradient_accumulation_under_no_sync(tmp_path):
    _check(tmp_path, n_accum=2)
