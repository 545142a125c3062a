"""Summarise an ncu --metrics gpu__time_duration.sum launch list: per-kernel share of the LAST forward."""
import csv, sys, re, collections
rows = [r for r in csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"'))]
names = [r["Kernel Name"] for r in rows]
last = max(i for i, n in enumerate(names) if "fbank_logmel_kernel" in n)
end = max(i for i, n in enumerate(names) if "skinny" in n or "small_linear_kernel" in n or "l2norm" in n) + 1
agg = collections.OrderedDict()
seq = []
for r in rows[last:end]:
    n = re.sub(r"\(.*", "", r["Kernel Name"]).replace("svb::", "")
    if "gemm_tc_kernel" in n:
        n += " grid=" + r["Grid Size"]
    t = float(r["Metric Value"]) / 1e3
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += t
    seq.append((n, t))
tot = sum(v[1] for v in agg.values())
print("launches in one forward: %d, total %.1f us" % (len(seq), tot))
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%7.1f us %5.1f%%  x%-3d %s" % (t, 100 * t / tot, c, n))
if len(sys.argv) > 2:
    for n, t in seq: print("   %8.1f  %s" % (t, n))
