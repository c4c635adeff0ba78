#!/usr/bin/env python
"""Summarises `ncu -i X.ncu-rep --page raw --csv` (one `--set full` capture) into a small CSV for profiles/:
per captured launch: duration, DRAM bytes, pipe utilisations, issue activity, top stall reasons."""
import csv
import sys

KEEP = [
    ("duration_us", "gpu__time_duration.sum"),
    ("grid", "launch__grid_size"),
    ("regs", "launch__registers_per_thread"),
    ("dram_read_MB", "dram__bytes_read.sum"),
    ("dram_write_MB", "dram__bytes_write.sum"),
    ("sm_throughput_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor_pipe_pct", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active"),
    ("xu_pipe_pct", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    ("fma_pipe_pct", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
    ("alu_pipe_pct", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
    ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("l2_throughput_pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
]


def to_float(v, unit, want):
    try:
        x = float(v.replace(",", ""))
    except ValueError:
        return v
    scale = {"nsecond": 1e-3, "ns": 1e-3, "usecond": 1.0, "us": 1.0, "msecond": 1e3, "ms": 1e3, "second": 1e6, "s": 1e6}
    if want == "duration_us":
        return round(x * scale.get(unit, 1.0), 2)
    if want.endswith("_MB"):
        return round(x * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1.0), 3)
    return round(x, 2)


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    stall_cols = [(h, i) for i, h in enumerate(hdr)
                  if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    out = csv.writer(sys.stdout)
    out.writerow(["kernel"] + [k for k, _ in KEEP] + ["top_stalls(per issue)"])
    for r in data:
        name = r[idx["Kernel Name"]].replace("void ", "").replace("unnamed>::", "")
        name = name.split("(")[0][:60]
        vals = []
        for k, m in KEEP:
            vals.append(to_float(r[idx[m]], units[idx[m]], k) if m in idx else "")
        st = []
        for h, i in stall_cols:
            try:
                st.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
            except ValueError:
                pass
        st.sort(reverse=True)
        out.writerow([name] + vals + [" ".join(f"{n}={v:.2f}" for v, n in st[:4])])


if __name__ == "__main__":
    main(sys.argv[1])
