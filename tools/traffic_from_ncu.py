"""DRAM traffic per launch of the dominant kernel from an `ncu --set full` capture exported with `--page raw --csv`:
mean of dram__bytes_read.sum + dram__bytes_write.sum over the launches whose kernel name matches.
usage: python tools/traffic_from_ncu.py raw.csv chain_kernel"""
import csv
import sys


def to_bytes(v, unit):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    h, units = rows[0], rows[1]
    k = h.index("Kernel Name"); r = h.index("dram__bytes_read.sum"); w = h.index("dram__bytes_write.sum")
    vals = [to_bytes(x[r], units[r]) + to_bytes(x[w], units[w]) for x in rows[2:] if sys.argv[2] in x[k]]
    print(len(vals), "launches, mean bytes", round(sum(vals) / len(vals)), "min", round(min(vals)), "max", round(max(vals)))


if __name__ == "__main__":
    main()
