"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (development aid)."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
agg = collections.OrderedDict()
for x in rows:
    k = x["Kernel Name"].replace("<unnamed>::", "").split("(")[0][:48]
    v = float(x["Metric Value"].replace(",", ""))
    u = x["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("%d launches, %.1f us total" % (len(rows), tot))
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-48s n=%4d total %10.1f us  avg %9.1f us  %5.1f%%" % (k, c, t, t / c, 100 * t / tot))
