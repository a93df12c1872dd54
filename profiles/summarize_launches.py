"""Summarise an `ncu --metrics gpu__time_duration.sum,... --csv` launch list per kernel."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
iK, iM, iV, iU = (hdr.index(n) for n in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[1:]:
    agg[r[iK].split("(")[0]][r[iM] + " [" + r[iU] + "]"].append(float(r[iV].replace(",", "")))
tot = sum(sum(m[k]) / len(m[k]) for m in agg.values() for k in m if k.startswith("gpu__time"))
for k, m in agg.items():
    line = ["%-24s" % k]
    for a, v in sorted(m.items()):
        line.append("%s avg %.1f (n=%d)" % (a, sum(v) / len(v), len(v)))
        if a.startswith("gpu__time"):
            line.append("share %.1f%%" % (100 * sum(v) / len(v) / tot))
    print("  ".join(line))
print("sum of per-kernel average durations: %.1f ns" % tot)
