"""Per-source-line stall samples / shared-memory wavefronts of one kernel from an ncu report (needs -lineinfo and
--import-source on).  usage: python scripts/ncu_src_lines.py report.ncu-rep ::regex:render_bwd:1 [top_n]   (ncu --kernel-id syntax)"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-id", kern],
                     capture_output=True, text=True).stdout
cur, hdr, agg = None, None, {}
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split('/')[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if r[0].isdigit() and hdr:
        d = dict(zip(hdr, r))

        def g(k):
            try:
                return int(d.get(k, '0').replace(',', ''))
            except ValueError:
                return 0
        agg[(cur, int(r[0]))] = (g('# Samples'), g('Instructions Executed'), g('L1 Wavefronts Shared'),
                                 g('L1 Wavefronts Shared Excessive'), g('stall_barrier'), g('stall_long_sb'),
                                 g('stall_short_sb'), g('stall_wait'), g('stall_math'), r[1][:80])
tot = [max(1, sum(v[i] for v in agg.values())) for i in range(9)]
print("kernel", kern, "totals: samples %d inst %d smem wavefronts %d (excess %d); stall samples barrier %d long_sb %d short_sb %d "
      "wait %d math %d" % tuple(tot))
print("--- top lines by stall samples (percent of the kernel's totals)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print("%-14s %4d samp %5.2f inst %5.2f wf %5.2f | bar %4.1f long %4.1f short %4.1f wait %4.1f math %4.1f | %s" % (
        k[0], k[1], 100 * v[0] / tot[0], 100 * v[1] / tot[1], 100 * v[2] / tot[2], 100 * v[4] / tot[4], 100 * v[5] / tot[5],
        100 * v[6] / tot[6], 100 * v[7] / tot[7], 100 * v[8] / tot[8], v[9]))
print("--- top lines by excessive shared-memory wavefronts")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][3])[:12]:
    print("%-14s %4d wf %5.2f%% excess %5.2f%% (%d of %d) | %s" % (k[0], k[1], 100 * v[2] / tot[2], 100 * v[3] / tot[3], v[3], v[2], v[9]))
