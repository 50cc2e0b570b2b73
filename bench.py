#!/usr/bin/env python
"""Benchmark of the renderer hot path (BASELINE.json metric: rendered rays/s, fwd+bwd, 128^2 images).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one forward+backward pass of the hot path over one synthetic batch: plane repack -> near/far ->
depth range -> fused coarse pass -> fine sampling -> fused fine pass -> fused backward -> gradient unpack.
Workload (config.workload): BASELINE.json configs[2] -- 128x128 rays per sample, 48+48 samples per ray, P=24
parts, fp32, GLOBAL batch 32 sample-sharded over the N ranks (32/N samples per GPU: strong scaling, the curve the
metric is quoted on).  ``--batch b`` instead gives every GPU b samples (weak scaling, the round-1 curve).
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
GLOBAL_BATCH = 32   # BASELINE.json configs[2]
SEED = 1236  # 1234 + config index 2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=None,
                    help="samples per GPU (weak scaling); default: configs[2]'s global batch 32 split over the ranks")
    ap.add_argument("--global-batch", type=int, default=GLOBAL_BATCH)
    ap.add_argument("--shard", default="balanced", choices=["balanced", "contiguous"],
                    help="how the global batch is dealt to the ranks (strong scaling)")
    ap.add_argument("--eager-e2e", action="store_true", help="end-to-end arm without the CUDA graph (GraphedRenderStep)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip timing the reference's eager path on this GPU")
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
def batch_part_frames(global_batch):
    """Scaled part frames of the whole synthetic batch (host data: 12 KB per sample)."""
    from enarf_gan_b200 import hostmath, synthetic as S
    per = [S.make_inputs(1, SIZE, seed=SEED + 17 * i) for i in range(global_batch)]
    ptc, bl = torch.cat([q["pose_to_camera"] for q in per]), torch.cat([q["bone_length"] for q in per])
    return hostmath.scaled_pose(hostmath.part_frames(ptc, bl, "center+head", S.SMPL_PARENTS)[0], 3.0)


def local_batch(args, rank, world):
    """Indices (into the global batch) of this rank's samples: configs[2]'s fixed global batch dealt to the ranks
    (strong scaling) -- by default balanced on a host-side cost proxy (rendering.balanced_shards), ``--shard contiguous``
    for plain blocks -- or ``--batch`` samples of its own on every rank (weak scaling)."""
    if args.batch is not None:
        return list(range(rank * args.batch, (rank + 1) * args.batch))
    from enarf_gan_b200 import rendering
    if args.shard == "balanced" and world > 1:
        return rendering.balanced_shards(rendering.sample_cost_proxy(batch_part_frames(args.global_batch)), world)[rank]
    lo, hi = rendering.shard_samples(args.global_batch, rank, world)
    return list(range(lo, hi))


def build_workload(samples, device, global_batch=None):
    """Synthetic batch resident in HBM: part frames, tri-planes (NCHW, as the reference produces them),
    per-sample decoder weights, rays, and fixed upstream gradients for the backward.  Sample g of the global batch is
    seeded by g alone, so the job renders the same 32 samples however many ranks share them."""
    from enarf_gan_b200 import hostmath, rendering, synthetic as S
    seed, B, first = SEED, len(samples), samples[0]
    per = [S.make_inputs(1, SIZE, seed=SEED + 17 * i) for i in samples]
    inp = {k: torch.cat([q[k] for q in per], dim=0) for k in per[0]}
    tp = torch.cat([S.make_triplane(1, num_part=PARTS, seed=SEED + 17 * i) for i in samples]).to(device)
    frames, bl = hostmath.part_frames(inp["pose_to_camera"], inp["bone_length"], "center+head", S.SMPL_PARENTS)
    pose = hostmath.scaled_pose(frames, 3.0).to(device)
    cframes, clen = hostmath.canonical_part_frames(S.canonical_pose(), S.SMPL_PARENTS, "center+head")
    cscale = (clen[None, :, None] / bl / 3.0)[:, :, 0].to(device)
    g = torch.Generator().manual_seed(seed + first)
    z = torch.randn(B, 256, generator=g)
    decoder = []
    for cin, cout in ((32, 64), (64, 64), (64, 4)):
        w = hostmath.modulated_weight(torch.randn(cout, cin, generator=g), torch.randn(cin, 256, generator=g),
                                      torch.ones(cin), z)
        decoder.append((w.to(device), (0.3 * torch.randn(cout, generator=g)).to(device)))
    n = SIZE * SIZE
    # Batch-global near / far (rendering.py:15-17) need the root depths of the WHOLE batch: every rank is handed the
    # full batch's part frames (12 KB per sample of host data) -- no collective on the render path
    pose_all = pose
    if global_batch is not None and global_batch != B:
        pose_all = batch_part_frames(global_batch).to(device)
    return dict(B=B, n=n, pose=pose, pose_all=pose_all, cscale=cscale, canon=cframes.to(device), tri_plane=tp, decoder=decoder,
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
    nearfar = ops.near_far(w["pose_all"])      # batch-global (rendering.py:15-17): the full batch's frames are local
    dmin, dmax, direction, valid = timed("depth_range", lambda: ops.depth_range(w["img"], w["inv_K"], w["pose"], nearfar))
    wb = timed("coarse", lambda: ops.render_coarse(scene, dmin, dmax, direction, NC))
    bins = timed("fine_sample", lambda: ops.fine_sample(wb, NF, seed=step_idx))
    color, mask, disp, fw, fd, pts, feat, pairs = timed("fine_fwd", lambda: ops._render_fine_fwd(
        scene, dmin, dmax, direction, bins, want_points=True, want_feat=True, want_pairs=True, skip_last=True))   # as a training step
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
            0.0, 0.0, ops.stream_ptr()))
    timed("fine_bwd", bwd)
    timed("unpack", lambda: ops.unpack_plane_grads(d_feat, d_tp, accumulate=False))
    w["last"] = (color, mask, d_tp, dw, bins, dmin, dmax, direction)
    w["ckpt_tile_base"] = pairs.tile_base
    return 8   # repack, near_far, depth_range, coarse, fine_sample, fine_fwd, fine_bwd, unpack


def mean_valid_parts(w):
    """v-bar of SURVEY.md 8(d): mean number of valid parts per point, for the coarse and the fine samples."""
    from enarf_gan_b200 import ops
    dbg = (lambda tag: (torch.cuda.synchronize(), print("[dbg] ok:", tag, file=sys.stderr))) if os.environ.get("ENARF_BENCH_DEBUG_SYNC") else (lambda tag: None)
    dbg("entry")
    scene = ops.SceneBuffers(w["pose"], w["cscale"], w["canon"], w["tri_plane"], w["decoder"])
    dbg("repack")
    _, _, _, _, bins, dmin, dmax, direction = w["last"]
    out = []
    cb = (torch.arange(NC, device=bins.device) + 0.5) / NC
    for u in (cb[None, None].expand(w["B"], w["n"], NC), bins):
        start, end = dmin[:, None] * direction, dmax[:, None] * direction
        pts = (start[..., None] * (1 - u[:, None]) + end[..., None] * u[:, None]).reshape(w["B"], 3, -1).contiguous()
        dbg("pts")
        _, _, m = ops.density_query(scene, pts, want_color=False, want_mask=True)
        dbg("query")
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


def build_e2e(w, device, graph=True):
    from enarf_gan_b200 import synthetic as S
    from enarf_gan_b200.generator import TriNARFGenerator
    cfg = dict(z_dim=256, background_ratio=0.7,
               nerf_params=dict(Nc=NC, Nf=NF, origin_location="center+head", coordinate_scale=3, render_bs=16384))
    gen = TriNARFGenerator(cfg, SIZE, num_bone=24, num_bone_param=23, parent_id=S.SMPL_PARENTS, black_background=True,
                           tri_plane_gen=_Planes(w["tri_plane"].clone()))
    gen.register_canonical_pose(S.canonical_pose())
    gen = gen.to(device).train()
    gen.nerf.rng_mode = "philox"
    from enarf_gan_b200 import ops
    gen.nerf.batch_near_far = ops.near_far(w["pose_all"])     # batch-global near / far without a collective
    pin = lambda t: t.contiguous().pin_memory()
    host = {k: pin(w["host"][k]) for k in ("pose_to_camera", "pose_to_world", "bone_length", "inv_intrinsics")}
    host["z"] = pin(S.make_latent(w["B"], seed=SEED))
    out_img = torch.empty(w["B"], 3, SIZE, SIZE).pin_memory()
    out_loss = torch.empty(1).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    d2h = out_img.numel() * 4 + 4
    if graph and w["B"] >= 2:
        # the public graphed step (enarf_gan_b200.graphed): generator forward + loss + backward captured once
        from enarf_gan_b200.graphed import GraphedRenderStep
        order = ("pose_to_camera", "pose_to_world", "bone_length", "z", "inv_intrinsics")
        gen = GraphedRenderStep(gen, e2e_loss, *[host[k] for k in order], device=device)
    return gen, host, out_img, out_loss, h2d, d2h


def e2e_loss(color, mask):
    return (color ** 2).mean() + ((1 - mask) ** 2).mean()


def step_e2e(gen, host, out_img, out_loss, device):
    if not isinstance(gen, torch.nn.Module):         # graphed: pinned host tensors in, static device tensors out
        color, mask, loss = gen(host["pose_to_camera"], host["pose_to_world"], host["bone_length"], host["z"],
                                host["inv_intrinsics"])
    else:
        dev = {k: v.to(device, non_blocking=True) for k, v in host.items()}
        gen.zero_grad(set_to_none=True)
        color, mask, fw, fd = gen(dev["pose_to_camera"], dev["pose_to_world"], dev["bone_length"], dev["z"],
                                  dev["inv_intrinsics"])
        loss = e2e_loss(color, mask)
        loss.backward()
    out_img.copy_(color.detach(), non_blocking=True)
    out_loss.copy_(loss.detach().reshape(1), non_blocking=True)


# ----------------------------------------------------------------------------------------------------
# Reference arms.  The UNMODIFIED reference (its eager grid_sample path) is imported through the test harness
# (tests/golden/ref_harness.py: /root/reference here, the staged archive oracle/_ref/reference_py.zip on the GPU box)
# and timed on the host cores (``cpu_baseline``, ``--impl reference``) and on this GPU (``gpu_reference``).  Only if
# the archive is missing does the CPU arm fall back to the oracle port.  None of this is on the product path.
# ----------------------------------------------------------------------------------------------------
def _paths():
    for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)


def reference_available():
    _paths()
    import ref_harness as H
    return H.available()


class ReferenceRun:
    """The reference's TriNARFGenerator on a synthetic batch of ``B`` samples at ``size`` x ``size`` rays."""

    def __init__(self, B, size, device, seed=1235, train=True):
        _paths()
        import ref_harness as H
        from enarf_gan_b200 import synthetic as S
        H.set_device(device)
        self.dev = torch.device(device)
        self.B, self.size = B, size
        gen = H.build_generator(size, S.canonical_pose(), S.SMPL_PARENTS, seed=seed, Nc=NC, Nf=NF)
        self.gen = gen.to(self.dev)
        self.gen.train(train)
        inp = S.make_inputs(B, size, seed=seed)
        self.inp = {k: v.to(self.dev) for k, v in inp.items()}
        self.tp = S.make_triplane(B, num_part=PARTS, seed=seed).to(self.dev).requires_grad_(train)
        self.z = S.make_latent(B, seed=seed).to(self.dev)
        nerf = self.gen.nerf
        orig = nerf._forward

        def forward_with_planes(*a, **k):       # models/narf.py:187-193: per-sample planes through model_input
            a = list(a)
            a[8]["tri_plane_feature"] = self.tp
            return orig(*a, **k)

        nerf._forward = forward_with_planes

    def step(self, backward=True):
        i = self.inp
        self.gen.zero_grad(set_to_none=True)
        self.tp.grad = None
        with torch.set_grad_enabled(backward):
            color, mask, fw, fd = self.gen(i["pose_to_camera"], i["pose_to_world"], i["bone_length"], self.z,
                                           i["inv_intrinsics"])
            if backward:
                loss = (color ** 2).mean() + ((1 - mask) ** 2).mean()
                loss.backward()
        return self.B * self.size * self.size


def cpu_reference(max_seconds=25.0, min_steps=3):
    """-> cpu_baseline dict: the reference's own CPU path, fwd+bwd, on ONE 128 x 128 sample of the workload
    (B = 1: a bounded sample; per-ray normalisation makes it comparable), all host threads, median of the steps; plus
    BASELINE.md section 2's figure: config 1 exactly (B=1, 64 x 64, forward only), 2 warm-up + 5 timed, median."""
    torch.set_num_threads(os.cpu_count() or 1)
    if not reference_available():
        return cpu_port()
    run = ReferenceRun(1, SIZE, "cpu", seed=1235)
    rays = run.step()                           # warm-up (allocator, thread pool)
    times = []
    while len(times) < min_steps or (sum(times) < max_seconds * 0.6 and len(times) < 9):
        t0 = time.perf_counter()
        run.step()
        times.append(time.perf_counter() - t0)
    med = sorted(times)[len(times) // 2]
    out = {"value": rays / med, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "reference",
           "sample": f"B=1, {SIZE}x{SIZE} rays (one sample of the workload), {NC}+{NF} samples/ray, P={PARTS}, fp32, forward+"
                     f"backward through the UNMODIFIED reference TriNARFGenerator (eager grid_sample path) on the host cores, "
                     f"median of {len(times)} steps after 1 warm-up", "seconds": sum(times), "ms_per_step": med * 1e3}
    c1 = ReferenceRun(1, 64, "cpu", seed=1235, train=False)
    for _ in range(2):
        c1.step(backward=False)
    t1 = []
    for _ in range(5):
        t0 = time.perf_counter()
        r1 = c1.step(backward=False)
        t1.append(time.perf_counter() - t0)
    m1 = sorted(t1)[2]
    out["config1_forward"] = {"value": r1 / m1, "unit": "rays/s", "points_per_s": r1 / m1 * (NC + NF), "ms": m1 * 1e3,
                              "what": "BASELINE configs[0] exactly: B=1, 64x64 rays, 48+48 samples, forward render, "
                                      "reference on CPU, 2 warm-up + 5 timed, median"}
    return out


def cpu_port():
    """Fallback when the reference archive is absent: the oracle port (torch-CPU restatement), same bounded sample."""
    _paths()
    import enarf_oracle as O
    from enarf_gan_b200 import synthetic as S
    B, size = 1, SIZE
    inp = S.make_inputs(B, size, seed=1235)
    tp = S.make_triplane(B, seed=1235)
    g = torch.Generator().manual_seed(1235)
    layers = [dict(mod_weight=torch.randn(cin, 256, generator=g), mod_bias=torch.ones(cin),
                   weight=torch.randn(cout, cin, generator=g), bias=0.3 * torch.randn(cout, generator=g))
              for cin, cout in ((32, 64), (64, 64), (64, 4))]
    sc = O.Scene(inp["pose_to_camera"], inp["bone_length"], inp["inv_intrinsics"], tp, layers,
                 torch.randn(B, 256, generator=g), S.canonical_pose(), S.SMPL_PARENTS)
    img = O.whole_image_grid(size, size, B)
    times = []
    for i in range(4):
        t0 = time.perf_counter()
        sc.tri_plane = sc.tri_plane.detach().requires_grad_()
        sc.eff = [(w.detach().requires_grad_(), b.detach().requires_grad_()) for w, b in sc.eff]
        out = O.render(sc, img, NC, NF, {"mode": "torch"})
        ((out["color"] ** 2).mean() + ((1 - out["mask"]) ** 2).mean()).backward()
        if i:
            times.append(time.perf_counter() - t0)
    med = sorted(times)[len(times) // 2]
    return {"value": size * size / med, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"B=1, {SIZE}x{SIZE} rays, {NC}+{NF} samples/ray, fwd+bwd through oracle/enarf_oracle.py (reference "
                      "archive oracle/_ref/reference_py.zip not staged)", "seconds": sum(times), "ms_per_step": med * 1e3}


def gpu_reference(device, B=2, steps=5, warmup=2):
    """the reference's eager PyTorch path on THIS B200, fwd+bwd, CUDA-event timed."""
    if not reference_available():
        return {"unavailable": "oracle/_ref/reference_py.zip not staged"}
    try:
        run = ReferenceRun(B, SIZE, str(device), seed=1235)
        for _ in range(warmup):
            run.step()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in ev:
            a.record()
            rays = run.step()
            b.record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)[steps // 2]
        peak = torch.cuda.max_memory_allocated() / 2 ** 30
        del run
        torch.cuda.empty_cache()
        return {"value": rays / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms, "batch": B, "peak_mem_gib": peak,
                "what": f"UNMODIFIED reference TriNARFGenerator.forward + backward (eager grid_sample path) on this GPU, "
                        f"B={B} x {SIZE}x{SIZE} rays, {NC}+{NF} samples, median of {steps} after {warmup} warm-up"}
    except Exception as e:  # noqa: BLE001  (a baseline must never take the bench line down)
        torch.cuda.empty_cache()
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on all host threads, same metric /
    config; every step is the bounded sample (one 128 x 128 sample of the batch)."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    if reference_available():
        run = ReferenceRun(1, SIZE, "cpu", seed=1235)
        kind, what = "reference", "UNMODIFIED reference TriNARFGenerator (eager grid_sample path) forward+backward on the host cores"
        step = run.step
    else:
        kind, what = "port", "oracle/enarf_oracle.py (reference archive not staged)"
        step = None
    if step is None:
        cb = cpu_port()
        value, ms, steps = cb["value"], cb["ms_per_step"], 3
    else:
        for _ in range(max(1, min(args.warmup, 2))):
            rays = step()
        times = []
        for _ in range(max(1, min(args.steps, 8))):
            t0 = time.perf_counter()
            rays = step()
            times.append(time.perf_counter() - t0)
        dt = sum(times) / len(times)
        value, ms, steps = rays / dt, dt * 1e3, len(times)
    line = base_line(args, value, ms, steps, world)
    sample = f"each step = one {SIZE}x{SIZE}-ray sample (B=1) of the workload, {NC}+{NF} samples/ray, fwd+bwd; {what}"
    line.update(impl="reference", gpu_launches=0,
                cpu_baseline={"value": value, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": kind, "sample": sample},
                e2e={"value": value, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    print(json.dumps(line))


def base_line(args, value, ms, steps, world):
    weak = args.batch is not None
    gb = args.batch * world if weak else args.global_batch
    return {"metric": "rendered_rays_per_s_fwd_bwd", "value": value, "unit": "rays/s", "n_gpus": world,
            "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": (f"BASELINE configs[2]: ENARF-GAN generator render fwd+bwd (tri-plane + MLP gradients), "
                                    f"{SIZE}x{SIZE} rays/sample, {NC}+{NF} samples/ray, P={PARTS} parts, global batch {gb} "
                                    + (f"= {args.batch} per GPU (weak)" if weak else f"sample-sharded over {world} GPU(s)")),
                       "global_batch": gb, "rays_per_step": gb * SIZE * SIZE,
                       "points_per_ray": NC + NF, "parallelism": f"sample-sharded x{world}, no data-path collective (every rank holds the batch's part frames)",
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
    samples = local_batch(args, rank, world)
    B = len(samples)
    assert B > 0, "more ranks than samples: use render_entire_img_sharded (ray blocks) for B < G"
    w = build_workload(samples, device, None if args.batch is not None else args.global_batch)
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
    overflow = int((w["ckpt_tile_base"] == -1).sum().item())     # tiles of the last step whose pairs did not fit the checkpoint
    # ---- end-to-end arm (host buffers in, image + loss out) ----
    gen, host, out_img, out_loss, h2d, d2h = build_e2e(w, device, graph=not args.eager_e2e)
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
    per_rank = None
    if world > 1:
        mine = torch.tensor([ms, ms_e2e, sum(kt.values())], device=device)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [[round(v, 4) for v in t.tolist()] for t in allr]
        ms, ms_e2e = max(r[0] for r in per_rank), max(r[1] for r in per_rank)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    gb = args.batch * world if args.batch is not None else args.global_batch
    rays = gb * SIZE * SIZE
    value = rays / (ms * 1e-3)
    line = base_line(args, value, ms, args.steps, world)
    line["clocks"] = clk
    line["gpu_launches"] = launches
    line["points_per_s"] = value * (NC + NF)
    line["e2e"] = {"value": rays / (ms_e2e * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "ms_per_step": ms_e2e, "api": ("TriNARFGenerator.forward + loss.backward (pinned host inputs)" if args.eager_e2e or B < 2 else
                           "graphed.GraphedRenderStep(TriNARFGenerator, loss): forward + loss + backward replayed from one CUDA "
                           "graph (pinned host inputs copied in, image + loss copied out)")}
    if per_rank is not None:
        # [device ms, e2e ms, sum of kernel ms] of every rank: the spread is the load imbalance (each rank's samples have
        # their own mean valid parts per point); there is no collective on the path
        line["per_rank_ms"] = per_rank
    line["samples_per_gpu"] = B
    line["config"]["sharding"] = "weak: own samples per rank" if args.batch is not None else \
        (f"{args.shard}" + (" (longest-processing-time-first on sum_parts 1/depth^2)" if args.shard == "balanced" else ""))
    line["pair_checkpoint_overflow_tiles"] = overflow
    # ---- roofline of the dominant kernel (rank 0's launches) ----
    (vc, ac), (vf, af) = mean_valid_parts(w)
    R = B * SIZE * SIZE
    # per (point, valid part) 396*s bytes, per ray the streamed I/O
    survey_bwd = 2 * R * NF * vf * 396 * 4 + R * (20 + NF * 4 + NF * 16)    # re-gather + reductions
    # what the backward really moves with the forward's checkpoints: per pair its 80-byte record and its
    # 128-byte feature line in, 12 x 128 B + 12 x 4 B of reductions out; per active point its 128-byte feature row;
    # per ray the streamed inputs
    ckpt_bwd = R * NF * vf * (80 + 128 + 396 * 4) + R * NF * af * 128 + R * (20 + NF * 4 + NF * 16)
    alg = {"coarse": R * NC * vc * 396 * 4 + R * (20 + NC * 4),
           "fine_fwd": R * NF * vf * 396 * 4 + R * (20 + NF * 4 + (NF - 1) * 4 + NF * 4 + NF * 16),
           "fine_bwd": survey_bwd}
    dom = max(alg, key=lambda k: kt[k])
    peak, peak_src = peaks()
    achieved = alg[dom] / (kt[dom] * 1e-3) / 1e9
    roof = {"bound": "l2-atomics + issue (NOT hbm: the planes are L2-resident, see dram_frac)", "kernel": dom,
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
            "frac_definition": "SURVEY 8(d) algorithmic bytes (396*s B per (point, part) pair, the backward counted as re-gather + "
                               "reductions) / CUDA-event kernel time / measured HBM copy peak.  These bytes are served by L2, so frac "
                               "is an L2-traffic rate expressed in HBM units, not DRAM utilisation (that is dram_frac); the checkpointed "
                               "backward no longer re-gathers, so what it really moves is moved_bytes_per_launch / moved_frac",
            "algorithmic_bytes_per_launch": alg[dom], "kernel_ms": kt[dom],
            "mean_valid_parts_per_point": {"coarse": vc, "fine": vf}, "active_point_fraction": {"coarse": ac, "fine": af},
            "kernel_ms_all": kt, "kernel_share_of_step": kt[dom] / ms, "traffic": None}
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic_r2.json")   # measured DRAM bytes per launch (ncu --set full)
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if dom in tj["kernels"] and tj.get("samples_per_gpu") == B:
            k = tj["kernels"][dom]
            roof["traffic"] = k["dram_bytes_read"] + k["dram_bytes_write"]
            roof["traffic_source"] = tj["source"]
            roof["dram_frac"] = roof["traffic"] / (kt[dom] * 1e-3) / 1e9 / peak
    if dom == "fine_bwd":
        roof["moved_bytes_per_launch"] = ckpt_bwd     # 208 B of checkpoint in + 1 584 B of reductions out per pair, 128 B per point
        roof["moved_frac"] = ckpt_bwd / (kt[dom] * 1e-3) / 1e9 / peak
        # the floor the backward cannot go under: its reductions at the measured L2-resident rates
        # (scripts/microbench/red_rate.cu, profiles/red_rate_r1.txt): 12 x 8 lane-level red.v4 per pair at 381.5 G/s and
        # 12 scattered 4-byte reductions per pair at 195.6 G/s
        pairs = R * NF * vf
        floor_ms = (pairs * 96 / 381.5e9 + pairs * 12 / 195.6e9) * 1e3
        roof["l2_red_floor_ms"] = floor_ms
        roof["l2_red_frac"] = floor_ms / kt[dom]
    line["roofline"] = roof
    if not args.no_gpu_reference and world == 1:
        line["gpu_reference"] = gpu_reference(device)
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_reference()
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
