"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches and total time per kernel."""
import collections, csv, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    k = r[ki][:70]; v = float(r[vi].replace(",", "")); u = r[ui]
    v = v / 1000.0 if u == "ns" else v * 1000.0 if u == "ms" else v
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
for k, (c, t) in agg.items():
    print(f"{c:5d} x {k:70s} {t:10.1f} us")
