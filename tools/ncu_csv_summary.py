#!/usr/bin/env python3
"""Summarise an `ncu --page raw --csv` export (one row per captured launch) into a text table under
profiles/: for every kernel name the LONGEST captured launch with duration, DRAM bytes, issue-active,
occupancy limits and the top stall reasons. Usage: python tools/ncu_csv_summary.py <raw.csv> <out.txt>"""
import csv
import sys

from ncu_summary import METRICS

EXTRA = ["launch__occupancy_limit_warps", "launch__occupancy_limit_blocks", "launch__waves_per_multiprocessor",
         "sm__warps_active.avg.per_cycle_active", "l1tex__t_sector_hit_rate.pct",
         "lts__t_bytes.sum", "smsp__inst_executed.sum"]


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return 0.0


def main():
    src, dst = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(src)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn, dur = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
    best = {}
    for d in data:
        name = d[kn].split("(")[0]
        if name not in best or num(d[dur]) > num(best[name][dur]):
            best[name] = d
    count = {}
    for d in data:
        count[d[kn].split("(")[0]] = count.get(d[kn].split("(")[0], 0) + 1
    with open(dst, "w") as out:
        out.write(f"# {src}: ncu --set full --clock-control none; {len(data)} launches captured, "
                  "the longest launch of every kernel shown\n")
        for name, d in sorted(best.items(), key=lambda kv: -num(kv[1][dur])):
            out.write(f"\n## {name[:110]}   (captured {count[name]}x)\n")
            for m in METRICS + [e for e in EXTRA if e not in METRICS]:
                if m in hdr:
                    i = hdr.index(m)
                    out.write(f"  {m:84s} {d[i]} {units[i]}\n")
    print("wrote", dst)


if __name__ == "__main__":
    sys.path.insert(0, __file__.rsplit("/", 1)[0])
    main()
