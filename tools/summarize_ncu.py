"""Per-kernel digest of an `ncu --set full` report: duration, DRAM traffic, pipe utilisation, occupancy, top stalls.

  ncu -i gpurun_out/r01h_full.ncu-rep --page raw --csv > /tmp/raw.csv
  python tools/summarize_ncu.py /tmp/raw.csv > profiles/r01h_ncu_full_summary.txt
  python tools/summarize_ncu.py /tmp/raw.csv --json > profiles/r01h_ncu_traffic.json   # per-launch DRAM bytes (bench.py reads it)
"""
import csv
import json
import re
import sys

WANT = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"), ("lts__t_bytes.sum", "l2_bytes"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu_pipe_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_pct")]
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1, "msecond": 1e3}


def short(n):
    return re.sub(r"<.*$", "", re.sub(r"\(.*$", "", re.sub(r"^void ", "", n))).split("::")[-1]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stall_cols = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    agg = {}
    for r in rows[2:]:
        name = short(r[idx["Kernel Name"]])
        a = agg.setdefault(name, {"n": 0, "vals": {}, "stalls": {}})
        a["n"] += 1
        for col, key in WANT:
            if col in idx and r[idx[col]] not in ("", "n/a"):
                try:
                    v = float(r[idx[col]].replace(",", "")) * UNIT.get(units[idx[col]], 1)
                except ValueError:
                    continue
                a["vals"][key] = a["vals"].get(key, 0.0) + v
        for col in stall_cols:
            try:
                a["stalls"][col[34:-23]] = a["stalls"].get(col[34:-23], 0.0) + float(r[idx[col]])
            except ValueError:
                pass
    if "--json" in sys.argv:
        print(json.dumps({k: {"launches": a["n"], "dram_bytes_per_launch": (a["vals"].get("dram_read", 0) + a["vals"].get("dram_write", 0)) / a["n"],
                              "us_per_launch_under_ncu": a["vals"].get("duration", 0) / a["n"]} for k, a in agg.items()}, indent=1))
        return
    print(f"# {sys.argv[1]}: per-launch averages (ncu --set full --clock-control none; times are cold-cache, serialised)")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["vals"].get("duration", 0)):
        n, v = a["n"], a["vals"]
        g = lambda key: v.get(key, 0.0) / n
        print(f"\n== {k}  ({n} launches captured)  grid {g('grid'):.0f} x {g('block'):.0f} threads, {g('regs'):.0f} regs")
        print(f"   duration {g('duration'):.2f} us | DRAM read {g('dram_read') / 1e6:.3f} MB write {g('dram_write') / 1e6:.3f} MB "
              f"({g('dram_pct'):.1f}% of peak) | L2 {g('l2_bytes') / 1e6:.2f} MB")
        print(f"   SM throughput {g('sm_pct'):.1f}% | issue active {g('issue_pct'):.1f}% | warps active {g('occupancy_pct'):.1f}% | "
              f"tensor pipe {g('tensor_pipe_pct'):.1f}% (tensor inst {g('tensor_inst'):.0f}) | XU pipe {g('xu_pipe_pct'):.1f}%")
        top = sorted(a["stalls"].items(), key=lambda kv: -kv[1])[:4]
        print("   top stalls (warps per issue): " + ", ".join(f"{s} {x / n:.2f}" for s, x in top))


if __name__ == "__main__":
    main()
