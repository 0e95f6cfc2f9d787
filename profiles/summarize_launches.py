"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total
device time and share.  usage: python profiles/summarize_launches.py <launches.csv>"""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
rows = list(csv.DictReader(lines))
agg = collections.OrderedDict()
for r in rows:
    k = r["Kernel Name"].split("(")[0][-70:]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += float(r["Metric Value"].replace(",", ""))
tot = sum(v[1] for v in agg.values())
print("total device time %.3f ms over %d launches" % (tot / 1e6, len(rows)))
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-72s n=%4d total=%10.1f us  share=%5.1f%%" % (k, v[0], v[1] / 1e3, 100 * v[1] / tot))
