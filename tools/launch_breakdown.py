"""Per-kernel totals of an ncu launch list (csv of gpu__time_duration.sum)."""
import csv, collections, sys
for f in sys.argv[1:]:
    rows = [r for r in csv.reader(open(f)) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    rows = rows[rows.index(hdr) + 1:]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    d = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for r in rows:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
        k = r[ki].split("(")[0][:48]
        d[k][0] += 1; d[k][1] += v; d[k][2] = max(d[k][2], v)
    tot = sum(v[1] for v in d.values())
    print(f"{f}: {sum(v[0] for v in d.values())} launches, {tot / 1e3:.3f} ms")
    for k, v in sorted(d.items(), key=lambda x: -x[1][1]):
        print(f"  {k:48s} {v[0]:5d} {v[1] / 1e3:9.3f} ms {100 * v[1] / tot:5.1f} %   max {v[2]:8.1f} us")
