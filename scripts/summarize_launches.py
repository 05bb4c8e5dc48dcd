"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (launches, total, share).

    python scripts/summarize_launches.py launches.csv [last_n_launches] > profiles/<name>.txt
"""
import collections
import csv
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[hi]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    data = [(r[ki], float(r[vi].replace(",", ""))) for r in rows[hi + 1:] if len(r) > vi]
    if len(sys.argv) > 2:
        data = data[-int(sys.argv[2]):]
    agg = collections.OrderedDict()
    for k, v in data:
        k = re.sub(r"^void ", "", k)
        k = re.sub(r"\(.*", "", k)[:120]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# {len(data)} launches, {tot / 1e3:.1f} us in kernels (cold-cache, serialised: compare SHARES)")
    print("# launches  total_us  share  us/launch  kernel")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{a[0]:6d} {a[1] / 1e3:12.1f} {100 * a[1] / tot:6.2f}% {a[1] / a[0] / 1e3:9.1f}  {k}")


if __name__ == "__main__":
    main()
