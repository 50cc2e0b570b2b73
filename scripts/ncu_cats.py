"""Instruction / stall-sample shares by code region for the fused kernels (development aid).
usage: ncu_cats.py src.csv   (src.csv = ncu --page source --csv --print-source cuda,sass)"""
import csv, sys
cur_file = cur_fn = None
agg = {}
for r in csv.reader(open(sys.argv[1])):
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split('/')[-1]; continue
    if r[0] == "Function Name": cur_fn = r[1][:44]; continue
    if r[0].isdigit() and len(r) > 8:
        try: ln = int(r[0]); inst = int(r[7]); samp = int(r[6])
        except ValueError: continue
        a = agg.setdefault((cur_fn, cur_file, ln), [0, 0]); a[0] += inst; a[1] += samp
def cat(fn, f, l):
    if f == 'tile.cuh':
        for hi, name in ((82, 'load decoder'), (123, 'phase0'), (183, 'phase1'), (206, 'prefix'), (224, 'chunk_end'),
                         (266, 'setup_pairs'), (312, 'gather'), (339, 'gemm_kvec'), (365, 'gemm_kmajor'), (9999, 'wgrad')):
            if l <= hi: return 'tile: ' + name
    if f == 'render.cu':
        for hi, name in ((67, 'layer epilogue'), (85, 'head'), (138, 'main/chunk loop'), (9999, 'composite/out')):
            if l <= hi: return 'render: ' + name
    if f == 'bwd.cu':
        for hi, name in ((151, 'flush/layer epi'), (205, 'scatter'), (245, 'tile load'), (300, 'composite'), (365, 'head'),
                         (395, 'G2 pass'), (9999, 'dgrad epilogues/loop')):
            if l <= hi: return 'bwd: ' + name
    if f == 'gather.cuh': return 'gather.cuh helpers'
    if f == 'common.cuh':
        if 41 <= l <= 50: return 'contract arith'
        if 76 <= l <= 95 or 144 <= l <= 153: return 'part tests'
        if 97 <= l <= 130: return 'bilinear+sigmoid'
        if 158 <= l <= 186: return 'ray_part_interval'
    return f
res = {}
for (fn, f, l), v in agg.items():
    a = res.setdefault(fn, {}).setdefault(cat(fn, f, l), [0, 0]); a[0] += v[0]; a[1] += v[1]
for fn, d in res.items():
    ti = sum(v[0] for v in d.values()); ts = sum(v[1] for v in d.values())
    print('==', fn, 'warp-inst', ti)
    for c, v in sorted(d.items(), key=lambda kv: -kv[1][1]):
        print('  %-30s inst %5.1f%%  samples %5.1f%%' % (c, 100 * v[0] / ti, 100 * v[1] / ts))
