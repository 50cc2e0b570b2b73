#!/usr/bin/env python
"""Benchmark of the renderer hot path (BASELINE.json metric: rendered rays/s, fwd+bwd, 128^2 images).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one forward+backward pass of the hot path over one synthetic batch: plane repack -> near/far ->
depth range -> fused coarse pass -> fine sampling -> fused fine pass -> fused backward -> gradient unpack.
Workload (config.workload): BASELINE.json configs[1]/[2] shapes -- 128x128 rays per sample, 48+48 samples per
ray, P=24 parts, fp32, batch 8 per GPU (weak scaling: global batch 8*N; N=4 is configs[2]'s batch 32).
Prints ONE JSON line on rank 0 (see the driver contract in the task statement).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

SIZE, NC, NF, PARTS = 128, 48, 48, 24
SEED = 1236  # 1234 + config index 2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="samples per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region
# ----------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------------
def build_workload(rank, B, device):
    """Synthetic batch resident in HBM: part frames, tri-planes (NCHW, as the reference produces them),
    per-sample decoder weights, rays, and fixed upstream gradients for the backward."""
    from enarf_gan_b200 import hostmath, rendering, synthetic as S
    seed = SEED + 1000 * rank
    inp = S.make_inputs(B, SIZE, seed=seed)
    tp = S.make_triplane(B, num_part=PARTS, seed=seed).to(device)
    frames, bl = hostmath.part_frames(inp["pose_to_camera"], inp["bone_length"], "center+head", S.SMPL_PARENTS)
    pose = hostmath.scaled_pose(frames, 3.0).to(device)
    cframes, clen = hostmath.canonical_part_frames(S.canonical_pose(), S.SMPL_PARENTS, "center+head")
    cscale = (clen[None, :, None] / bl / 3.0)[:, :, 0].to(device)
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(B, 256, generator=g)
    decoder = []
    for cin, cout in ((32, 64), (64, 64), (64, 4)):
        w = hostmath.modulated_weight(torch.randn(cout, cin, generator=g), torch.randn(cin, 256, generator=g),
                                      torch.ones(cin), z)
        decoder.append((w.to(device), (0.3 * torch.randn(cout, generator=g)).to(device)))
    n = SIZE * SIZE
    return dict(B=B, n=n, pose=pose, cscale=cscale, canon=cframes.to(device), tri_plane=tp, decoder=decoder,
                inv_K=inp["inv_intrinsics"].to(device), img=rendering.pixel_grid(SIZE, SIZE, B, device),
                g_color=torch.randn(B, 3, n, generator=g).to(device), g_mask=torch.randn(B, n, generator=g).to(device),
                host=inp)


class KernelTimer:
    """CUDA events around individual launches on the launching (current) stream."""

    def __init__(self):
        self.spans = {}

    def span(self, name):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.spans.setdefault(name, []).append((a, b))
        return a, b

    def totals(self):
        return {k: sum(a.elapsed_time(b) for a, b in v) / len(v) for k, v in self.spans.items()}


def step_device(w, step_idx, timer=None, world=None):
    """One fwd+bwd pass with every input already in HBM.  Returns the launch count."""
    from enarf_gan_b200 import ops

    def timed(name, fn):
        if timer is None:
            return fn()
        a, b = timer.span(name)
        a.record()
        out = fn()
        b.record()
        return out

    B, n = w["B"], w["n"]
    dev = w["pose"].device
    scene = timed("repack", lambda: ops.SceneBuffers(w["pose"], w["cscale"], w["canon"], w["tri_plane"], w["decoder"]))
    pose_all = w["pose"]
    if world is not None:          # batch-global near/far (rendering.py:15-17) across the sharded batch
        import torch.distributed as dist
        gathered = [torch.empty_like(w["pose"]) for _ in range(world)]
        dist.all_gather(gathered, w["pose"])
        pose_all = torch.cat(gathered, dim=0)
    nearfar = ops.near_far(pose_all)
    dmin, dmax, direction, valid = timed("depth_range", lambda: ops.depth_range(w["img"], w["inv_K"], w["pose"], nearfar))
    wb = timed("coarse", lambda: ops.render_coarse(scene, dmin, dmax, direction, NC))
    bins = timed("fine_sample", lambda: ops.fine_sample(wb, NF, seed=step_idx))
    color, mask, disp, fw, fd, pts, feat, pairs = timed("fine_fwd", lambda: ops._render_fine_fwd(
        scene, dmin, dmax, direction, bins, want_points=True, want_feat=True, want_pairs=True))
    d_tp = torch.empty_like(w["tri_plane"])     # as ops.RenderFine.backward: feature channels are stored by unpack
    for b in range(B):                          # per sample: contiguous slices -> vectorised fills
        d_tp[b, 96:].zero_()
    d_feat = torch.zeros(B, 3, scene.H, scene.W, 32, device=dev)
    dw = [torch.zeros_like(t) for t in scene.w]
    db = [torch.zeros_like(t) for t in scene.b]
    p = ops.ptr

    def bwd():
        ops.check(ops.lib().enarf_render_fine_bwd(
            scene.ref, p(dmin), p(dmax), p(direction), p(bins), p(pts), p(feat), pairs.ref, n, NF, p(w["g_color"]), p(w["g_mask"]),
            None, None,
            p(d_feat), d_tp[:, 96:].data_ptr(), d_tp.stride(0), p(dw[0]), p(dw[1]), p(dw[2]), p(db[0]), p(db[1]), p(db[2]),
            ops.stream_ptr()))
    timed("fine_bwd", bwd)
    timed("unpack", lambda: ops.unpack_plane_grads(d_feat, d_tp, accumulate=False))
    w["last"] = (color, mask, d_tp, dw, bins, dmin, dmax, direction)
    return 8   # repack, near_far, depth_range, coarse, fine_sample, fine_fwd, fine_bwd, unpack


def mean_valid_parts(w):
    """v-bar of SURVEY.md 8(d): mean number of valid parts per point, for the coarse and the fine samples."""
    from enarf_gan_b200 import ops
    scene = ops.SceneBuffers(w["pose"], w["cscale"], w["canon"], w["tri_plane"], w["decoder"])
    _, _, _, _, bins, dmin, dmax, direction = w["last"]
    out = []
    cb = (torch.arange(NC, device=bins.device) + 0.5) / NC
    for u in (cb[None, None].expand(w["B"], w["n"], NC), bins):
        start, end = dmin[:, None] * direction, dmax[:, None] * direction
        pts = (start[..., None] * (1 - u[:, None]) + end[..., None] * u[:, None]).reshape(w["B"], 3, -1).contiguous()
        _, _, m = ops.density_query(scene, pts, want_color=False, want_mask=True)
        cnt = torch.zeros_like(m)
        for p in range(PARTS):
            cnt += (m >> p) & 1
        out.append((cnt.float().mean().item(), (cnt > 0).float().mean().item()))
    return out


# ----------------------------------------------------------------------------------------------------
# end-to-end arm: the public API (TriNARFGenerator.forward + backward) fed from pinned HOST buffers
# ----------------------------------------------------------------------------------------------------
class _Planes(torch.nn.Module):
    """Stands in for the tri-plane producer (StyleGAN2-ADA generator, out of scope): one learnable plane set
    per sample, resident on the device like any network activation."""

    def __init__(self, tp):
        super().__init__()
        self.tp = torch.nn.Parameter(tp)

    def forward(self, z, bone_length, truncation_psi=1):
        return self.tp if z.shape[0] == self.tp.shape[0] else self.tp[:z.shape[0]]


def build_e2e(w, device):
    from enarf_gan_b200 import synthetic as S
    from enarf_gan_b200.generator import TriNARFGenerator
    cfg = dict(z_dim=256, background_ratio=0.7,
               nerf_params=dict(Nc=NC, Nf=NF, origin_location="center+head", coordinate_scale=3, render_bs=16384))
    gen = TriNARFGenerator(cfg, SIZE, num_bone=24, num_bone_param=23, parent_id=S.SMPL_PARENTS, black_background=True,
                           tri_plane_gen=_Planes(w["tri_plane"].clone()))
    gen.register_canonical_pose(S.canonical_pose())
    gen = gen.to(device).train()
    gen.nerf.rng_mode = "philox"
    pin = lambda t: t.contiguous().pin_memory()
    host = {k: pin(w["host"][k]) for k in ("pose_to_camera", "pose_to_world", "bone_length", "inv_intrinsics")}
    host["z"] = pin(S.make_latent(w["B"], seed=SEED))
    out_img = torch.empty(w["B"], 3, SIZE, SIZE).pin_memory()
    out_loss = torch.empty(1).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    d2h = out_img.numel() * 4 + 4
    return gen, host, out_img, out_loss, h2d, d2h


def step_e2e(gen, host, out_img, out_loss, device):
    dev = {k: v.to(device, non_blocking=True) for k, v in host.items()}
    gen.zero_grad(set_to_none=True)
    color, mask, fw, fd = gen(dev["pose_to_camera"], dev["pose_to_world"], dev["bone_length"], dev["z"],
                              dev["inv_intrinsics"])
    loss = (color ** 2).mean() + ((1 - mask) ** 2).mean()
    loss.backward()
    out_img.copy_(color.detach(), non_blocking=True)
    out_loss.copy_(loss.detach().reshape(1), non_blocking=True)


# ----------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path, timed on the host cores (bounded sample)
# ----------------------------------------------------------------------------------------------------
CPU_SAMPLE = "B=1, 128x128 rays (one sample of the config-2 batch), 48+48 samples/ray, P=24, fp32, forward+backward through oracle/enarf_oracle.py"


def cpu_step(state=None):
    """One fwd+bwd render of the bounded sample with the CPU oracle.  Returns (state, rays, seconds)."""
    for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import enarf_oracle as O
    from enarf_gan_b200 import synthetic as S
    torch.set_num_threads(os.cpu_count() or 1)
    if state is None:
        B, size = 1, 128
        inp = S.make_inputs(B, size, seed=1235)
        tp = S.make_triplane(B, seed=1235)
        g = torch.Generator().manual_seed(1235)
        layers = [dict(mod_weight=torch.randn(cin, 256, generator=g), mod_bias=torch.ones(cin),
                       weight=torch.randn(cout, cin, generator=g), bias=0.3 * torch.randn(cout, generator=g))
                  for cin, cout in ((32, 64), (64, 64), (64, 4))]
        sc = O.Scene(inp["pose_to_camera"], inp["bone_length"], inp["inv_intrinsics"], tp, layers,
                     torch.randn(B, 256, generator=g), S.canonical_pose(), S.SMPL_PARENTS)
        state = (sc, O.whole_image_grid(size, size, B), size)
    sc, img, size = state
    t0 = time.perf_counter()
    sc.tri_plane = sc.tri_plane.detach().requires_grad_()
    sc.eff = [(w.detach().requires_grad_(), b.detach().requires_grad_()) for w, b in sc.eff]
    out = O.render(sc, img, NC, NF, {"mode": "torch"})
    loss = (out["color"] ** 2).mean() + ((1 - out["mask"]) ** 2).mean()
    loss.backward()
    return state, size * size, time.perf_counter() - t0


def run_reference(args, rank, world):
    """--impl reference: the CPU port of the reference path on all host threads, same metric / config."""
    if rank != 0:
        return
    state = None
    for _ in range(max(1, min(args.warmup, 1))):
        state, rays, _ = cpu_step(state)
    times = []
    for _ in range(max(1, min(args.steps, 5))):
        state, rays, dt = cpu_step(state)
        times.append(dt)
    dt = sum(times) / len(times)
    value = rays / dt
    line = base_line(args, value, dt * 1e3, len(times))
    line.update(impl="reference", gpu_launches=0,
                cpu_baseline={"value": value, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
                              "sample": CPU_SAMPLE},
                e2e={"value": value, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    line["config"]["note"] = ("/root/reference cannot travel to the GPU box; the arm times oracle/enarf_oracle.py, the "
                              "torch-CPU restatement pinned against the reference's own outputs")
    print(json.dumps(line))


def base_line(args, value, ms, steps):
    return {"metric": "rendered_rays_per_s_fwd_bwd", "value": value, "unit": "rays/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"ENARF-GAN generator render fwd+bwd, {SIZE}x{SIZE} rays/sample, {NC}+{NF} samples/ray, "
                                   f"P={PARTS} parts, batch {args.batch} per GPU (BASELINE configs[1]/[2] shapes)",
                       "global_batch": args.batch * max(1, args.gpus), "rays_per_step": args.batch * max(1, args.gpus) * SIZE * SIZE,
                       "points_per_ray": NC + NF, "parallelism": f"sample-sharded x{max(1, args.gpus)}, no data-path collective",
                       "l2": "inputs larger than L2 (tri-planes 44 MB/sample x batch, re-packed every step)"}}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback for the product path)"
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    from enarf_gan_b200 import _abi
    _abi.lib()
    w = build_workload(rank, args.batch, device)
    wd = world if world > 1 else None

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident arm ----
    for i in range(max(3, args.warmup)):
        step_device(w, i, world=wd)
    sync()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    timer = KernelTimer()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    t0.record()
    for i in range(args.steps):
        launches += step_device(w, 100 + i, timer, world=wd)
    t1.record()
    sync()
    ms = t0.elapsed_time(t1) / args.steps
    kt = timer.totals()
    # ---- end-to-end arm (host buffers in, image + loss out) ----
    gen, host, out_img, out_loss, h2d, d2h = build_e2e(w, device)
    for _ in range(max(3, args.warmup)):
        step_e2e(gen, host, out_img, out_loss, device)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_e2e(gen, host, out_img, out_loss, device)
    e1.record()
    sync()
    ms_e2e = e0.elapsed_time(e1) / args.steps
    clk = clocks.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms, ms_e2e], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = t.tolist()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    rays = args.batch * world * SIZE * SIZE
    value = rays / (ms * 1e-3)
    line = base_line(args, value, ms, args.steps)
    line["n_gpus"] = world
    line["clocks"] = clk
    line["gpu_launches"] = launches
    line["points_per_s"] = value * (NC + NF)
    line["e2e"] = {"value": rays / (ms_e2e * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "ms_per_step": ms_e2e, "api": "TriNARFGenerator.forward + loss.backward (pinned host inputs)"}
    # ---- roofline of the dominant kernel (rank 0's launches) ----
    (vc, ac), (vf, af) = mean_valid_parts(w)
    R = args.batch * SIZE * SIZE
    alg = {"coarse": R * NC * vc * 396 * 4 + R * (20 + NC * 4),
           "fine_fwd": R * NF * vf * 396 * 4 + R * (20 + NF * 4 + (NF - 1) * 4 + NF * 4 + NF * 16),
           "fine_bwd": 2 * R * NF * vf * 396 * 4 + R * (20 + NF * 4 + NF * 16)}
    dom = max(alg, key=lambda k: kt[k])
    peak, peak_src = peaks()
    achieved = alg[dom] / (kt[dom] * 1e-3) / 1e9
    traffic = traffic_src = None     # measured DRAM bytes per launch of that kernel, from the committed ncu capture
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic_r1.json")
    if os.path.exists(tpath) and args.batch == 8:
        tj = json.load(open(tpath))
        if dom in tj["kernels"]:
            traffic = tj["kernels"][dom]["dram_bytes_read"] + tj["kernels"][dom]["dram_bytes_write"]
            traffic_src = tj["source"]
    line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": alg[dom], "kernel_ms": kt[dom],
                        "mean_valid_parts_per_point": {"coarse": vc, "fine": vf},
                        "active_point_fraction": {"coarse": ac, "fine": af},
                        "kernel_ms_all": kt, "kernel_share_of_step": kt[dom] / ms}
    if not args.no_cpu_baseline:
        state, crays, _ = cpu_step(None)   # warm-up (allocator, thread pool)
        cts = []
        while len(cts) < 3 and sum(cts) < 20.0:   # about 10 s of CPU work on the box's 16 threads
            state, crays, dt = cpu_step(state)
            cts.append(dt)
        dt = sum(cts) / len(cts)
        line["cpu_baseline"] = {"value": crays / dt, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
                                "sample": CPU_SAMPLE + f", mean of {len(cts)} steps", "seconds": sum(cts)}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
