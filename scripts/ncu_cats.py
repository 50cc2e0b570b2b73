"""Instruction / stall-sample shares by code region for the fused kernels (development aid).
usage: ncu_cats.py src.csv   (src.csv = ncu --page source --csv --print-source cuda,sass, one kernel)
Regions are found from marker strings in the current sources, not hard-coded line numbers."""
import csv, os, sys
ROOT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "enarf-gan_b200", "csrc")
MARK = {
    "tile.cuh": [("load_decoder_smem(", "load decoder"), ("void tile_points(", "tile: phase 0/1"),
                 ("exclusive prefix of pairs-per-point", "tile: prefix"), ("int chunk_end(", "chunk_end"),
                 ("void setup_pairs(", "setup_pairs (cta)"), ("void gather_chunk(", "gather_chunk (cta)"),
                 ("register-tiled FP32 GEMMs", "ffma helpers"), ("void gather32_warp(", "warp: sub-batch bookkeeping"),
                 ("setup: one lane per pair", "warp: pair setup"), ("the four groups take points", "warp: gather")],
    "render.cu": [("render_kernel(", "render: tile prologue"), ("chunks of 32 active points", "render: chunk loop + epilogue"),
                  ("---- outputs", "render: outputs/composite")],
    "bwd.cu": [("wgrad_zero(", "bwd: wgrad flush"), ("void decoder_layer(", "bwd: layer epilogue"),
               ("void scatter_chunk(", "bwd: scatter"), ("render_bwd_kernel(", "bwd: tile load"),
               ("compositing backward, one warp", "bwd: composite"), ("per chunk: records", "bwd: chunk prologue / X load"),
               ("head forward + backward", "bwd: head"), ("G2 = (W3^T g3)", "bwd: G2 pass"),
               ("dW2 += G2^T A1", "bwd: wgrad/dgrad calls + epilogues")],
    "common.cuh": [("float cmul(", "contract arith"), ("void load_part_params(", "load part params"),
                   ("bool part_local(", "part transforms"), ("struct Bilin", "bilinear setup/fetch + sigmoid"),
                   ("uint4 philox4x32_10(", "philox"), ("bool part_valid(", "part_valid"),
                   ("bool ray_part_interval(", "ray_part_interval")],
    "gather.cuh": [("float lrelu_s2(", "lrelu"), ("float unit_bin(", "unit_bin"), ("float4 load_feat4(", "texel load (LDG + address)"),
                   ("float4 gather_point(", "gather_point")],
    "mlp_mma.cuh": [("load_decoder_mma(", "load decoder"), ("void mma_tf32(", "mma + split helpers"),
                    ("void decode32_mma(", "decoder L1"), ("layer-1 activations -> the warp", "decoder A1 store"), ("---- layer 2: 64 -> 64", "decoder L2+L3")],
    "mma_smem.cuh": [("struct FragA", "bwd mma: fragment loads + split"), ("void mma3_tiles(", "bwd mma: mma issue"),
                     ("void gemm_pts_mma(", "bwd mma: gemm loop"), ("void wgrad_mma(", "bwd mma: wgrad loop")],
}
bounds = {}
for f, marks in MARK.items():
    lines = open(os.path.join(ROOT, f)).read().split("\n")
    b = []
    for key, name in marks:
        hit = [i + 1 for i, l in enumerate(lines) if key in l]
        if hit: b.append((hit[0], name))
    bounds[f] = sorted(b)
def cat(f, l):
    name = f
    for lo, n in bounds.get(f, []):
        if l >= lo: name = f.split(".")[0] + ": " + n if not n.startswith(("tile", "warp", "render", "bwd")) else n
    return name
cur_file = cur_fn = None
agg = {}
for r in csv.reader(open(sys.argv[1])):
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split('/')[-1]; continue
    if r[0] == "Function Name": cur_fn = r[1][:44]; continue
    if r[0].isdigit() and len(r) > 8:
        try: ln = int(r[0]); inst = int(r[7]); samp = int(r[6])
        except ValueError: continue
        a = agg.setdefault((cur_fn, cat(cur_file, ln)), [0, 0]); a[0] += inst; a[1] += samp
for fn in sorted(set(k[0] for k in agg)):
    d = {k[1]: v for k, v in agg.items() if k[0] == fn}
    ti = sum(v[0] for v in d.values()) or 1; ts = sum(v[1] for v in d.values()) or 1
    print('==', fn, 'warp-inst', ti, 'samples', ts)
    for c, v in sorted(d.items(), key=lambda kv: -kv[1][1]):
        print('  %-44s inst %5.1f%%  samples %5.1f%%' % (c, 100 * v[0] / ti, 100 * v[1] / ts))
