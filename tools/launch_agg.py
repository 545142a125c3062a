"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name (any workload)."""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    k = r[4][:90]
    unit, val = r[-2], float(r[-1].replace(",", ""))
    us = val / 1000.0 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1000.0)
    agg.setdefault(k, [0, 0.0])
    agg[k][0] += 1
    agg[k][1] += us
tot = sum(t for _, t in agg.values())
print("total %.1f us over %d launches" % (tot, len(rows)))
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 16]:
    print("%9.1f us %5.1f%%  x%-4d %s" % (t, 100 * t / tot, n, k))
