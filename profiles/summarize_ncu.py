"""Key counters of an `ncu --set full` report, one block per captured launch.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > /tmp/raw.csv
    python profiles/summarize_ncu.py /tmp/raw.csv > profiles/<name>.txt
"""
import csv
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "sm__cycles_elapsed.max",
]


def main(path):
    rows = list(csv.reader(open(path)))
    header, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(header)}
    for r in rows[2:]:
        print("==", r[col["Kernel Name"]][:100])
        for w in WANT:
            if w in col:
                print(f"   {w:66s} {r[col[w]]:>16s} {units[col[w]]}")


if __name__ == "__main__":
    main(sys.argv[1])
