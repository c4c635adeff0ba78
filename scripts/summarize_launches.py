#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share."""
import collections
import csv
import re
import sys


def main(path, top=45):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        name = re.sub(r"^void ", "", re.sub(r"\(.*", "", row["Kernel Name"]))
        name = name.replace("afb::<unnamed>::", "afb::")
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
    ours = sum(t for k, (n, t) in agg.items() if k.startswith("afb::"))
    print(f"# {path}: {sum(a[0] for a in agg.values())} launches, {tot:.1f} us serialised device time; "
          f"adaface_b200 kernels {ours:.1f} us ({100 * ours / tot:.1f} %)")
    print("total_us,share_pct,launches,avg_us,kernel")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{t:.1f},{100 * t / tot:.1f},{n},{t / n:.1f},{k[:120]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)
