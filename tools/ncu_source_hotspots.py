"""Source lines of one kernel ranked by executed instructions / stall samples, from an `ncu --set full --import-source on` report:

  python tools/ncu_source_hotspots.py gpurun_out/r02_full.ncu-rep rt_metrics 32 [min_share]

`units` scales the instruction counts (e.g. CTAs or regions of the launch): the first column is warp instructions per unit."""
import csv
import subprocess
import sys

rep, kern, units = sys.argv[1], sys.argv[2], float(sys.argv[3])
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.015
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "-k", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if len(r) > 3 and r[0] == "Line No"][0]
hdr = rows[hi]
ix, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
res, tot, ts = [], 0, 0
for r in rows[hi + 1:]:
    if len(r) <= ix or r[0] == "":
        continue
    try:
        n, s = int(r[ix]), int(r[ismp])
    except ValueError:
        continue
    tot += n
    ts += s
    res.append((n, s, r[0], r[1]))
print(f"{kern}: {tot} warp instructions ({tot / units:.0f} per unit), {ts} stall samples")
for n, s, line, src in res:
    if s > ts * thr or n > tot * thr:
        print(f"{n / units:9.1f} {100 * s / max(ts, 1):5.1f}% {line:>5s} {src[:130]}")
