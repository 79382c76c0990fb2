"""Summarises an ncu --csv launch list (gpu__time_duration.sum) per kernel name."""
import collections, csv, io, sys
txt = open(sys.argv[1]).read()
r = csv.DictReader(io.StringIO(txt[txt.index('"ID"'):]))
agg = collections.OrderedDict(); n = 0
for row in r:
    if row.get("Metric Name") != "gpu__time_duration.sum": continue
    name = row["Kernel Name"][:90]; v = float(row["Metric Value"].replace(",", "")); u = row["Metric Unit"]
    v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v; n += 1
tot = sum(a[1] for a in agg.values())
print("%d launches, %.1f us total" % (n, tot))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1])[: int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print("%10.1f us %4dx %5.1f%%  %s" % (a[1], a[0], 100 * a[1] / tot, k))
