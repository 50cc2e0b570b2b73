"""CUDA-graph capture of the generator's render step ("CUDA-graph the whole generator render").

A render step is ~10 kernels of ours plus ~150 small torch launches (modulated weights, losses, fills, autograd
bookkeeping); eager, the host side costs ~1 ms per step whatever the batch -- 2 % of a 32-sample step, 15 % of the 4
samples a GPU renders when configs[2] is sharded over 8.  ``GraphedRenderStep`` captures generator forward + loss +
backward once (static shapes, B >= 2: the B == 1 path compacts its valid rays on the host) and replays it:

    step = GraphedRenderStep(gen, loss_fn, pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics)
    color, mask, loss = step(pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics)   # parameter .grad filled
    optimizer.step()

Inputs are copied into the graph's static tensors (host tensors are welcome: the copies are the step's H2D traffic);
outputs are the graph's static tensors (valid until the next call).  Fresh fine-pass draws on every replay come from a
Philox seed that lives on the device (``nerf.rng_seed_dev``) and is bumped inside the graph.
"""
import torch


class GraphedRenderStep:
    def __init__(self, gen, loss_fn, pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics, warmup=3, device=None):
        nerf = gen.nerf
        dev = device or next(gen.parameters()).device
        if pose_to_camera.shape[0] < 2:
            raise ValueError("GraphedRenderStep needs B >= 2: the B == 1 render compacts its valid rays through the host")
        if getattr(nerf, "rng_mode", "torch") != "philox":
            raise ValueError("GraphedRenderStep draws in-kernel: set gen.nerf.rng_mode = 'philox'")
        self.gen, self.loss_fn = gen, loss_fn
        self.static = [t.detach().to(dev, copy=True) for t in (pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics)]
        if getattr(nerf, "rng_seed_dev", None) is None:
            nerf.rng_seed_dev = torch.zeros(1, device=dev, dtype=torch.int64)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                   # warm-up off the capture stream (lazy module loads, autotuning)
            for _ in range(warmup):
                gen.zero_grad(set_to_none=True)
                self._step()
        torch.cuda.current_stream(dev).wait_stream(side)
        gen.zero_grad(set_to_none=True)
        self.graph = torch.cuda.CUDAGraph()
        # (the warm-up's AccumulateGrad nodes lived on the side stream: the mismatch warning is expected and harmless)
        warn = torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch
        warn(False)
        try:
            with torch.cuda.graph(self.graph):
                self.color, self.mask, self.loss = self._step()
        finally:
            warn(True)

    def _step(self):
        color, mask, _, _ = self.gen(*self.static)
        loss = self.loss_fn(color, mask)
        loss.backward()
        return color.detach(), mask.detach(), loss.detach()

    def __call__(self, pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics):
        for dst, src in zip(self.static, (pose_to_camera, pose_to_world, bone_length, z, inv_intrinsics)):
            dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.color, self.mask, self.loss
