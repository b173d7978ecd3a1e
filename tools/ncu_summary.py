"""Prints selected metrics of every kernel in an .ncu-rep (run in the build container)."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum", "smsp__inst_executed.sum", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [i for i, h in enumerate(hdr) if h in WANT]
    ki = hdr.index("Kernel Name")
    print("| kernel | " + " | ".join(f"{hdr[i]} [{units[i]}]" for i in idx) + " |")
    print("|---|" + "---:|" * len(idx))
    for r in rows[2:]:
        print(f"| `{r[ki].split('(')[0].replace('void ', '')}` | " + " | ".join(r[i] for i in idx) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
