"""Summarise an ncu --csv launch list (gpu__time_duration.sum) per kernel name."""
import csv, sys, collections
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", "")); u = r[ui]
    ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
    a = agg.setdefault(r[ki][:60], [0, 0.0]); a[0] += 1; a[1] += ms
tot = sum(a[1] for a in agg.values())
for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:60s} n={n:5d} total={ms:9.3f} ms avg={1e3*ms/n:8.1f} us share={100*ms/tot:5.1f}%")
print(f"total {tot:.2f} ms over {sum(a[0] for a in agg.values())} launches")
