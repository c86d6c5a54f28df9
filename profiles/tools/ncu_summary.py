#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the per-launch table kept under profiles/.

    ncu -i gpurun_out/.../capture.ncu-rep --page raw --csv > raw.csv
    python profiles/tools/ncu_summary.py raw.csv > profiles/rN_<what>_summary.txt
"""
import csv
import sys

COLS = [
    ("time ms", "gpu__time_duration.sum", 1.0),
    ("DRAM rd GB", "dram__bytes_read.sum", 1.0),
    ("DRAM wr GB", "dram__bytes_write.sum", 1.0),
    ("L2->L1 GB", "l1tex__m_xbar2l1tex_read_bytes.sum", 1.0),
    ("L1TEX %", "l1tex__throughput.avg.pct_of_peak_sustained_active", 1.0),
    ("L2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("warps %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    ("IPC", "sm__inst_executed.avg.per_cycle_elapsed", 1.0),
    ("Minst", "smsp__inst_executed.sum", 1e-6),
    ("regs", "launch__registers_per_thread", 1.0),
    ("grid", "launch__grid_size", 1.0),
    ("block", "launch__block_size", 1.0),
    ("L1 hit %", "l1tex__t_sector_hit_rate.pct", 1.0),
    ("L2 hit %", "lts__t_sector_hit_rate.pct", 1.0),
    ("st long-sb", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", 1.0),
    ("st barrier", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", 1.0),
    ("st short-sb", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", 1.0),
]


def to_unit(v, unit, want):
    """ncu prints bytes / times in varying units; normalise to GB / ms."""
    x = float(v.replace(",", ""))
    u = unit.lower()
    if want == "GB":
        return x * {"byte": 1e-9, "kbyte": 1e-6, "mbyte": 1e-3, "gbyte": 1.0}.get(u, 1.0)
    if want == "ms":
        return x * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)
    return x


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    print("kernel".ljust(64) + "".join(name.rjust(12) for name, _, _ in COLS))
    for r in data:
        name = r[idx["Kernel Name"]]
        name = name.replace("void ", "").split("(")[0][:62]
        out = name.ljust(64)
        for label, key, scale in COLS:
            if key not in idx or r[idx[key]] == "":
                out += "-".rjust(12)
                continue
            want = "GB" if "GB" in label else ("ms" if label == "time ms" else "")
            out += f"{to_unit(r[idx[key]], units[idx[key]], want) * scale:12.3f}"
        print(out)


if __name__ == "__main__":
    main(sys.argv[1])
