"""Summarise an ncu report (`ncu --set full ... -o prof`) as one line per kernel launch.

    python bench/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/ncu_step_kernels_r2.txt

Runs `ncu -i <rep> --page raw --csv` (works without a GPU) and prints duration, tensor-pipe share, DRAM bytes,
L2 throughput, achieved occupancy, registers, grid and cluster size of every captured launch, in launch order.
"""
import csv
import io
import subprocess
import sys

COLS = [("us", "gpu__time_duration.sum"), ("tensor%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        ("rdMB", "dram__bytes_read.sum"), ("wrMB", "dram__bytes_write.sum"),
        ("l2%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("occ%", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("regs", "launch__registers_per_thread"),
        ("grid", "launch__grid_size"), ("cl", "launch__cluster_dim_x")]
SCALE = {"nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6, "ns": 1e-3, "us": 1.0, "ms": 1e3,
         "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def main(rep: str) -> int:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(head)}
    name_col = idx.get("Kernel Name", idx.get("Function Name"))
    print("%-68s" % "kernel" + "".join("%10s" % c for c, _ in COLS))
    for r in body:
        vals = []
        for _, m in COLS:
            if m not in idx or r[idx[m]] in ("", "n/a"):
                vals.append(float("nan"))
                continue
            v = float(r[idx[m]].replace(",", ""))
            vals.append(v * SCALE.get(units[idx[m]], 1.0))
        name = r[name_col]
        name = name.replace("dtf::", "").split("(")[0]
        print("%-68s" % name[:67] + "".join("%10.1f" % v for v in vals))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1]))
