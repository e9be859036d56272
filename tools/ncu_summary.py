"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel totals for this library's kernels.
usage: python tools/ncu_summary.py launches.csv [out.csv]"""
import collections, csv, sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
agg = collections.OrderedDict()
other = [0, 0.0]
for row in csv.DictReader(lines):
    name = row["Kernel Name"]
    ns = float(row["Metric Value"].replace(",", ""))
    if row["Metric Unit"] in ("us", "usecond"):
        ns *= 1e3
    if name.startswith("pg_") or "pg::" in name:
        short = name.split("(")[0].replace("void ", "")
        a = agg.setdefault(short, [0, 0.0])
        a[0] += 1
        a[1] += ns
    else:
        other[0] += 1
        other[1] += ns
tot = sum(a[1] for a in agg.values())
out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout
out.write("kernel,launches,total_us,avg_us,share_of_library_kernels\n")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.write("%s,%d,%.1f,%.2f,%.4f\n" % (k, a[0], a[1] / 1e3, a[1] / 1e3 / a[0], a[1] / tot))
out.write("(torch kernels: input generation and self-check outside the timed region),%d,%.1f,,\n" % (other[0], other[1] / 1e3))
