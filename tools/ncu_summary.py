#!/usr/bin/env python
"""`ncu -i X.ncu-rep --page raw --csv --print-units base` (stdin or file)  ->  a small tracked CSV:
one row per (kernel, metric), averaged over the kernel's launches in the report (`launches` = how many), followed
by one row per individual launch (kernel name suffixed `#i`, launches = 1) so that a warm-up launch or an outlier
stays visible.  bench.py reads `roofline.traffic` from such a file (profiles/r02_ncu_mlp_up_batch2.csv; it matches
kernel names by substring and takes the first, i.e. the averaged, row).
usage: ncu_summary.py raw.csv out.csv ["comment line"]"""
import csv
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_tensor.sum", "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__registers_per_thread",
           "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct"]


def main():
    raw, out = sys.argv[1], sys.argv[2]
    comment = sys.argv[3] if len(sys.argv) > 3 else ""
    with open(raw) as f:
        rows = [r for r in csv.reader(l for l in f if not l.startswith("=="))]
    hdr = rows[0]
    units = rows[1]
    ki = hdr.index("Kernel Name")
    acc = {}
    each = []
    seen = {}
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        idx = seen[r[ki]] = seen.get(r[ki], -1) + 1
        for m in METRICS:
            if m in hdr:
                j = hdr.index(m)
                try:
                    v = float(r[j].replace(",", ""))
                except ValueError:
                    continue
                a = acc.setdefault((r[ki], m, units[j]), [0.0, 0])
                a[0] += v
                a[1] += 1
                each.append((f"{r[ki]}#{idx}", m, units[j], v))
    with open(out, "w", newline="") as f:
        if comment:
            f.write("# " + comment + "\n")
        w = csv.writer(f)
        w.writerow(["kernel", "metric", "unit", "value", "launches"])
        for (k, m, u), (s, n) in acc.items():
            w.writerow([k, m, u, s / n, n])
        multi = {k for (k, _, _), (_, n) in acc.items() if n > 1}
        for k, m, u, v in each:
            if k.rsplit("#", 1)[0] in multi:
                w.writerow([k, m, u, v, 1])


if __name__ == "__main__":
    main()
