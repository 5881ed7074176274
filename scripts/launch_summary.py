"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total, share."""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    v = v / 1e3 if r[ui] in ("ns", "nsecond") else v
    a = agg.setdefault(r[ki].split("(")[0][:70], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("%-72s %5s %12s %7s" % ("kernel", "n", "total us", "share"))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-72s %5d %12.1f %6.1f%%" % (k, a[0], a[1], 100 * a[1] / tot))
print("%-72s %5s %12.1f" % ("ALL", "", tot))
