"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.

    python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/<name>_summary.txt
"""
import collections
import csv
import sys


def main(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("void ", "")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) * scale[r[ui]]
    tot = sum(a[1] for a in agg.values())
    print(f"# {path}: {len(rows) - 1} launches, {tot:.3f} ms of kernel time (cold-cache, serialised under ncu)")
    print(f"{'kernel':48s} {'launches':>8s} {'total ms':>10s} {'avg us':>10s} {'share':>7s}")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k[:48]:48s} {a[0]:8d} {a[1]:10.3f} {a[1] / a[0] * 1e3:10.1f} {a[1] / tot:7.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
