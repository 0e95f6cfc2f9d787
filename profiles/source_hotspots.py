"""Per-CUDA-source-line instruction counts from an ncu report:
  ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > src.csv
  python profiles/source_hotspots.py src.csv [topN]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = collections.OrderedDict()
fname, hdr = "", None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        line, src = r[0], r[1]
        i_inst = hdr.index("Instructions Executed")
        i_thr = hdr.index("Thread Instructions Executed")
        i_smp = hdr.index("# Samples")
        try:
            inst, thr, smp = float(r[i_inst] or 0), float(r[i_thr] or 0), float(r[i_smp] or 0)
        except ValueError:
            continue
        a = agg.setdefault((fname, line), [src, 0.0, 0.0, 0.0])
        if src:
            a[0] = src
        a[1] += inst; a[2] += thr; a[3] += smp
tot = sum(a[1] for a in agg.values()) or 1.0
tots = sum(a[3] for a in agg.values()) or 1.0
print("total warp instructions %.3g, samples %d" % (tot, tots))
for (f, line), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-16s:%4s inst %5.1f%%  samples %5.1f%%  thr/inst %4.1f | %s" % (
        f, line, 100 * a[1] / tot, 100 * a[3] / tots, a[2] / a[1] if a[1] else 0, a[0].strip()[:100]))
