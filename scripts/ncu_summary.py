#!/usr/bin/env python
"""Text summaries of the `ncu --set full` captures in gpurun_out/ (scripts/gpu_ncu.sh) for profiles/.

usage: python scripts/ncu_summary.py TAG      -> profiles/TAG_ncu_<kernel>.txt, prints the DRAM bytes per launch
Reads the reports with `ncu -i ... --page details --csv` and `--page raw --csv`; no GPU needed.
"""
import csv
import io
import subprocess
import sys

KEEP = ["Duration", "DRAM Throughput", "Memory Throughput", "Max Bandwidth", "Mem Busy", "Compute (SM) Throughput",
        "Executed Ipc Active", "Issue Slots Busy", "L1/TEX Hit Rate", "L2 Hit Rate", "Registers Per Thread",
        "Dynamic Shared Memory Per Block", "Static Shared Memory Per Block", "Block Limit Registers",
        "Block Limit Shared Mem", "Theoretical Occupancy", "Achieved Occupancy", "Grid Size", "Block Size",
        "Shared Memory Configuration Size"]
RAW = ["dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "gpu__time_duration.sum",
       "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
       "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
       "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
       "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    tag = sys.argv[1]
    for k in ("sort_scatter", "linear_forward", "linear_update", "sort_hist"):
        rep = f"gpurun_out/r02_{k}.ncu-rep"
        rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "details", "--csv"]))))
        hdr = rows[0]
        iid, ik, im, iv, iu = (hdr.index(x) for x in ("ID", "Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
        out = [f"# ncu --set full --clock-control none --import-source on, bench.py --steps 1 --warmup 1 (default workload), "
               f"launches 10-12 of the kernel; `ncu -i {rep} --page details --csv` filtered, then selected raw metrics"]
        for r in rows[1:]:
            if len(r) > iv and r[im] in KEEP:
                out.append("\t".join([r[iid], r[ik][:70], r[im], r[iv], r[iu]]))
        raw = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
        rh = raw[0]
        tot = []
        for r in raw[2:]:
            d = dict(zip(rh, r))
            out.append("\t".join([d.get("ID", ""), d.get("Kernel Name", "")[:70]] + [f"{m}={d.get(m)}" for m in RAW]))
            # (units of dram__bytes are in the units row: Mbyte here)
            tot.append((d.get("Kernel Name", "")[:60], float(d["dram__bytes_read.sum"]) + float(d["dram__bytes_write.sum"]),
                        float(d["gpu__time_duration.sum"])))
        units = dict(zip(rh, raw[1]))
        out.append(f"# units: dram__bytes_read.sum [{units.get('dram__bytes_read.sum')}], gpu__time_duration.sum [{units.get('gpu__time_duration.sum')}]")
        open(f"profiles/{tag}_ncu_{k}.txt", "w").write("\n".join(out) + "\n")
        print(k, [(n, round(b, 1), round(t, 1)) for n, b, t in tot], "mean dram", round(sum(b for _, b, _ in tot) / len(tot), 1),
              units.get("dram__bytes_read.sum"))


if __name__ == "__main__":
    main()
