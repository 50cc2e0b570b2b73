"""How many of the 128-B feature-plane lines that one tile's (point, part) pairs touch are shared between pairs of the
SAME tile?  (Would shared-memory privatisation of the backward's scatter have anything to merge?)  Read from the
forward's own pair checkpoint at the bench workload (B = 8, 128 x 128, 48+48).   python scripts/dup_texels.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from enarf_gan_b200 import ops

dev = torch.device("cuda", 0)
w = bench.build_workload(list(range(8)), dev)
scene = ops.SceneBuffers(w["pose"], w["cscale"], w["canon"], w["tri_plane"], w["decoder"])
dmin, dmax, direction, _ = ops.depth_range(w["img"], w["inv_K"], w["pose"])
wb = ops.render_coarse(scene, dmin, dmax, direction, 48)
bins = ops.fine_sample(wb, 48, seed=1)
out = ops._render_fine_fwd(scene, dmin, dmax, direction, bins, want_points=True, want_feat=True, want_pairs=True)
pairs = out[-1]
torch.cuda.synchronize()
n_pairs = int(pairs.cursor)
rec = pairs.rec[:n_pairs]
base = pairs.tile_base.long()
base = torch.where(base < 0, torch.full_like(base, n_pairs), base)    # -1 not checkpointed / -2 empty tile: sorted last, own no pair
tiles = base.numel()
# tile id of every pair: tiles were handed consecutive slices in the order they claimed them
order = torch.argsort(base)
starts = base[order]
tile_of_pair = torch.empty(n_pairs, dtype=torch.long, device=dev)
tile_rank = torch.bucketize(torch.arange(n_pairs, device=dev), starts, right=True) - 1
tile_of_pair = order[tile_rank]
W = scene.W
off = rec[:, :3].contiguous().view(torch.int32).long() // 32                  # texel index of (y0, x0) per plane
lines = torch.stack([off, off + 1, off + W, off + W + 1], dim=2).reshape(n_pairs, 12)
key = tile_of_pair[:, None] * (1 << 32) + lines
uniq = torch.unique(key.reshape(-1))
per_tile_total = torch.bincount(tile_of_pair, minlength=tiles) * 12
per_tile_uniq = torch.bincount((uniq >> 32), minlength=tiles)
nz = per_tile_total > 0
dup = 1.0 - per_tile_uniq[nz].double() / per_tile_total[nz].double()
print(f"{n_pairs} pairs in {int(nz.sum())} non-empty tiles of {tiles}; lines touched {int(per_tile_total.sum())}, "
      f"distinct within their tile {int(per_tile_uniq.sum())}")
print(f"share of line updates that hit a line already touched by the same tile: mean {dup.mean():.3f}, "
      f"median {dup.median():.3f}, p90 {dup.quantile(0.9):.3f}, max {dup.max():.3f}")
hist = torch.histc(dup.float(), bins=10, min=0, max=1)
print("histogram over tiles (duplicate share 0.0-0.1, ..., 0.9-1.0):", [int(v) for v in hist.tolist()])
# the same question per 64-point chunk is bounded by the per-tile figure; across the whole sample:
k2 = (tile_of_pair // (tiles // 8))[:, None] * (1 << 32) + lines
print(f"distinct lines per sample / line updates per sample: {torch.unique(k2.reshape(-1)).numel() / lines.numel():.3f}")
# merge potential at finer granularity: windows of consecutive pairs in list order (what one warp of the backward's
# scatter could merge in registers: pairs of consecutive samples of a ray)
tex = off.reshape(n_pairs, 3)            # texel index of the footprint per plane
for win in (4, 8, 12, 20, 40, 95):
    m = (n_pairs // win) * win
    t = tex[:m].reshape(-1, win, 3)
    wid = torch.arange(t.shape[0], device=dev)[:, None, None].expand_as(t)
    k3 = (wid * (1 << 28) + t).reshape(-1)
    print(f"window of {win:3d} consecutive pairs: distinct footprints / footprints = {torch.unique(k3).numel() / k3.numel():.3f}")
# the same per 128-B line (a footprint is 4 lines; shifted footprints share 2)
for win in (12, 24, 95):
    m = (n_pairs // win) * win
    t = lines[:m].reshape(-1, win * 12)
    wid = torch.arange(t.shape[0], device=dev)[:, None].expand_as(t)
    k3 = (wid * (1 << 28) + t).reshape(-1)
    print(f"window of {win:3d} consecutive pairs: distinct lines / line updates = {torch.unique(k3).numel() / k3.numel():.3f}")
