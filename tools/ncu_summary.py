#!/usr/bin/env python
"""Selected metrics of every launch in `ncu --set full` reports, one CSV line per (kernel, shape), last launch kept.

    python tools/ncu_summary.py out.csv "label 1" a.ncu-rep ["label 2" b.ncu-rep ...]

Reads each report with `ncu -i ... --page raw --csv` (run where ncu is installed; no GPU needed).
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "sm__cycles_active.avg", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__cycles_active.avg",
]


def main():
    out, pairs = sys.argv[1], sys.argv[2:]
    lines, header = [], None
    for label, rep in zip(pairs[0::2], pairs[1::2]):
        text = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(text)))
        names, units = rows[0], rows[1]
        idx = {m: names.index(m) for m in METRICS if m in names}
        kcol = names.index("Kernel Name")
        if header is None:
            header = ["shape", "Kernel Name"] + [f"{m} [{units[idx[m]]}]" if units[idx[m]] else m for m in METRICS if m in idx]
        last = {}
        for r in rows[2:]:
            last[r[kcol]] = r                       # later launches overwrite earlier ones: warm instruction caches
        for k, r in last.items():
            lines.append([label, k] + [r[idx[m]] for m in METRICS if m in idx])
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(header)
        w.writerows(lines)
    print(f"{len(lines)} kernels -> {out}")


if __name__ == "__main__":
    main()
