"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total ms, share."""
import collections
import csv
import sys


def summarise(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "second": 1e3}
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        a = agg[r[ki]]
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    tot = sum(v[1] for v in agg.values())
    out = [f"# {path}: {len(rows) - 1} launches, {tot:.3f} ms of kernel time (cold-cache, serialised by ncu)",
           f"# {'total ms':>10} {'n':>5} {'share':>6} {'ms/launch':>10}  kernel"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"  {v[1]:10.3f} {v[0]:5d} {100 * v[1] / tot:5.1f}% {v[1] / v[0]:10.4f}  {k[:140]}")
    return "\n".join(out)


if __name__ == "__main__":
    print(summarise(sys.argv[1]))
