"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel count, time, share."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None
agg = collections.OrderedDict()
for r in rows:
    if r and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        try:
            v = float(d["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        k = d["Kernel Name"].split("(")[0][-60:]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
tot = sum(v[1] for v in agg.values())
print("%-62s %6s %12s %7s %10s" % ("kernel", "count", "total us", "share", "avg us"))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-62s %6d %12.1f %6.1f%% %10.1f" % (k, v[0], v[1] / 1e3, 100 * v[1] / tot, v[1] / 1e3 / v[0]))
print("total %.1f us over %d launches" % (tot / 1e3, sum(v[0] for v in agg.values())))
