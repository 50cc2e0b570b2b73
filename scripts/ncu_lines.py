"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by source line (development aid).
usage: python scripts/ncu_lines.py src.csv [top_n]"""
import csv, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
cur_file = cur_fn = None
agg = {}
for r in csv.reader(open(path)):
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split('/')[-1]; continue
    if r[0] == "Function Name": cur_fn = r[1][:44]; continue
    if r[0] == "Line No": continue
    if r[0].isdigit() and len(r) > 8:
        try:
            ln = int(r[0]); inst = int(r[7]); samp = int(r[6])
        except ValueError:
            continue
        a = agg.setdefault((cur_fn, cur_file, ln), [0, 0, r[1][:100]])
        a[0] += inst; a[1] += samp
for fn in sorted(set(k[0] for k in agg)):
    tot = sum(v[0] for k, v in agg.items() if k[0] == fn); ts = sum(v[1] for k, v in agg.items() if k[0] == fn)
    print("=====", fn, "total inst", tot, "samples", ts)
    items = sorted(((v[0], v[1], k[1], k[2], v[2]) for k, v in agg.items() if k[0] == fn), reverse=True)
    for it in items[:topn]:
        print("%5.1f%% inst %5.1f%% samp %s:%d  %s" % (100 * it[0] / tot, 100 * it[1] / max(ts, 1), it[2], it[3], it[4]))
