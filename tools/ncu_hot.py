"""Top stall sites of one kernel from `ncu -i X.ncu-rep --page source --csv` (SASS view)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = next(r for r in rows if r and r[0] == "Address")
idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows if len(r) == len(hdr) and r[0].startswith("0x")]
S = idx["# Samples"]
tot = sum(int(r[S]) for r in data)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(int(r[idx[h]] or 0) for r in data) for h in stalls}
print("total samples", tot, "| by reason:", ", ".join("%s %.0f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(data, key=lambda r: -int(r[S]))[:n]:
    st = sorted(((h[6:], int(r[idx[h]] or 0)) for h in stalls), key=lambda kv: -kv[1])[:2]
    print("%5.1f%%  %-66s %s" % (100.0 * int(r[S]) / max(tot, 1), r[idx["Source"]].strip()[:66], st))
