"""Compact per-launch summary of an `ncu --set full` capture exported with `--page raw --csv`.
usage: python tools/ncu_summary.py raw.csv > profiles/ncu_rNN_<kernel>_summary.txt"""
import csv
import sys

WANT = [
    ("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"), ("launch__shared_mem_per_block_dynamic", "dyn smem"),
    ("gpu__time_duration.sum", "duration"), ("sm__cycles_elapsed.max", "cycles elapsed"), ("sm__cycles_active.avg", "SM cycles active (avg)"),
    ("sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe (imma) % of peak while active"),
    ("sm__inst_executed_pipe_tensor_subpipe_imma.avg.pct_of_peak_sustained_active", "imma issue % of peak"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "TMEM cycles active % elapsed"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("smsp__inst_executed.sum", "warp instructions"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        for key, label in WANT:
            if key in h:
                i = h.index(key)
                print(f"{label:45s} {r[i][:90]} {units[i]}")
        print("-" * 100)


if __name__ == "__main__":
    main()
