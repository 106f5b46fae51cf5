"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list.

  python tools/summarize_launches.py gpurun_out/r01b_launches.csv [> profiles/r01b_launch_summary.txt]

Times under ncu are cold-cache and serialised: read the SHARE column, not the absolute numbers.
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = re.sub(r"<.*$", "", name)
    return name.split("::")[-1]


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        rows.append((short(r["Kernel Name"]), us, r["Grid Size"], r["Block Size"]))
    agg = OrderedDict()
    for n, us, g, b in rows:
        a = agg.setdefault(n, [0, 0.0, g, b])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print(f"# {path}: {len(rows)} launches, {tot:.1f} us total under ncu (cold cache, serialised)")
    print(f"{'kernel':44s} {'launches':>8s} {'total_us':>10s} {'avg_us':>8s} {'share':>7s}  grid / block (first launch)")
    for n, (c, us, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n[:44]:44s} {c:8d} {us:10.1f} {us / c:8.2f} {100 * us / tot:6.1f}%  {g} / {b}")


if __name__ == "__main__":
    main(sys.argv[1])
