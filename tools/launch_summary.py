"""Summarise an ncu launch list (csv of gpu__time_duration.sum) per step: python tools/launch_summary.py file.csv STEPS"""
import collections
import csv
import re
import sys


def main(path, steps, top=40, seq=False):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    gi = hdr.index("Grid Size")
    data = rows[1:]
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[data[0][ui]]
    agg, tot = collections.OrderedDict(), 0.0
    for r in data:
        k = re.sub(r"\(.*", "", r[ki])[:90]
        v = float(r[vi].replace(",", "")) * scale / steps
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1.0 / steps
        a[1] += v
        tot += v
    print(f"# {path}: {len(data) / steps:.0f} launches/step, {tot:.3f} ms/step (cold-cache, serialised)")
    for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{v:8.3f} ms {100 * v / tot:5.1f}% {c:6.1f}  {k}")
    if seq:
        per = len(data) // steps
        for i, r in enumerate(data[-per:]):
            print(i, f"{float(r[vi].replace(',', '')) * scale * 1e3:8.1f}us", re.sub(r"void ", "", r[ki])[:120], r[gi])


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 2, seq="--seq" in sys.argv)
