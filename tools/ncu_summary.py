#!/usr/bin/env python3
"""Summarise one gpu_round.sh visit into profiles/: launch list (per-kernel share) and the
`ncu --set full` raw page (selected metrics per captured launch).
Usage: python tools/ncu_summary.py <tag> [gpurun_out]"""
import collections, csv, subprocess, sys, os

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]


def main():
    tag = sys.argv[1]
    src = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out"
    out = open(os.path.join("profiles", f"{tag}_ncu_summary.txt"), "w")
    lp = os.path.join(src, f"{tag}_launches.csv")
    if os.path.exists(lp):
        rows = list(csv.DictReader(l for l in open(lp) if l.startswith('"')))
        agg = collections.OrderedDict()
        for r in rows:
            k = r["Kernel Name"].split("(")[0][:72]
            a = agg.setdefault(k, [0, 0.0])
            a[0] += 1
            a[1] += float(r["Metric Value"].replace(",", ""))
        tot = sum(v[1] for v in agg.values())
        out.write(f"# launch list ({lp}): ncu --metrics gpu__time_duration.sum --clock-control none\n")
        out.write(f"# {len(rows)} launches, {tot/1e6:.3f} ms summed (cold-cache, serialised; includes the bench's set-up renders)\n")
        for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
            out.write(f"{k:74s} n={v[0]:4d} {v[1]/1e3:10.1f} us share={v[1]/tot:6.3f}\n")
    rp = os.path.join(src, f"{tag}_prof.ncu-rep")
    if os.path.exists(rp):
        raw = subprocess.run(["ncu", "-i", rp, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units, data = rows[0], rows[1], rows[2:]
        out.write(f"\n# full capture ({rp}): ncu --set full --clock-control none --import-source on\n")
        # per-launch DRAM traffic of the largest launch of every kernel -> bench.py roofline.traffic
        names = {"preprocess_fwd_kernel": "preprocess_fwd", "duplicate_tiles_kernel": "duplicate_keys",
                 "radix_onesweep_kernel": "radix_onesweep", "blend_fwd_kernel": "blend_fwd",
                 "blend_bwd_kernel": "blend_bwd", "geom_bwd_kernel": "geom_bwd",
                 "blend_fwd2_kernel": "blend_fwd", "blend_bwd2_kernel": "blend_bwd"}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        traffic = {}
        for d in data:
            kn = d[hdr.index("Kernel Name")]
            for pat, short in names.items():
                if pat in kn:
                    b = float(d[ir].replace(",", "")) * scale[units[ir]] + float(d[iw].replace(",", "")) * scale[units[iw]]
                    traffic[short] = max(traffic.get(short, 0.0), b)
        import json
        with open(os.path.join("profiles", f"{tag}_traffic.json"), "w") as tf:
            json.dump({"source": f"ncu --set full capture {tag}_prof.ncu-rep (1 view of c3), dram__bytes_read.sum + "
                                 "dram__bytes_write.sum per launch", "bytes_per_launch": traffic}, tf, indent=1)
        for d in data:
            out.write(f"\n## {d[hdr.index('Kernel Name')][:100]}\n")
            for m in METRICS:
                if m in hdr:
                    i = hdr.index(m)
                    out.write(f"  {m:84s} {d[i]} {units[i]}\n")
    out.close()
    print("wrote", out.name)


if __name__ == "__main__":
    main()
