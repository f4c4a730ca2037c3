"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python tools/launch_summary.py launches.csv"""
import collections, csv, re, sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
r = csv.reader(lines)
hdr = next(r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg, tot = collections.defaultdict(lambda: [0, 0.0]), 0.0
for row in r:
    v = float(row[vi].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(row[ui], 1e-6)
    name = re.sub(r"^void |<unnamed>::|\(.*", "", row[ki])
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
print("total %.3f ms over %d launches (cold-cache, serialised under ncu: shares, not absolutes)" % (tot, sum(a[0] for a in agg.values())))
print("| kernel | launches | ms | share | avg us |\n|---|---|---|---|---|")
for k, (n, ms) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("| %s | %d | %.3f | %.2f%% | %.1f |" % (k[:70], n, ms, 100 * ms / tot, 1e3 * ms / n))
