"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv).  usage: launch_table.py file.csv [first last]"""
import collections, csv, io, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = [r for r in csv.DictReader(io.StringIO("".join(lines))) if r.get("Metric Name") == "gpu__time_duration.sum"]
if len(sys.argv) > 3: rows = rows[int(sys.argv[2]):int(sys.argv[3])]
agg = collections.OrderedDict()
for r in rows:
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v = v / 1e3 if u.startswith("n") else (v * 1e3 if u.startswith("m") else v)
    k = r["Kernel Name"].split("(")[0][:60]
    a = agg.setdefault(k, [0, 0.0, []]); a[0] += 1; a[1] += v; a[2].append(v)
tot = sum(a[1] for a in agg.values())
for k, a in agg.items():
    print("%-60s n=%4d total=%10.1f us (%5.1f %%)  last=%s" % (k, a[0], a[1], 100 * a[1] / tot, ["%.0f" % x for x in a[2][-10:]]))
print("total %.1f us over %d launches" % (tot, len(rows)))
