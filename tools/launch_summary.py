"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel name.
usage: python tools/launch_summary.py gpurun_out/launches.csv [skip_first_n_launches]"""
import collections
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
    h = rows[hi]; k = h.index('Kernel Name'); v = h.index('Metric Value'); u = h.index('Metric Unit')
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for r in rows[hi + 1 + skip:]:
        if len(r) <= v:
            continue
        try:
            t = float(r[v].replace(',', ''))
        except ValueError:
            continue
        if r[u] == 'us':
            t *= 1e3
        elif r[u] == 'ms':
            t *= 1e6
        n = r[k].split('(')[0][:70]
        agg[n][0] += 1; agg[n][1] += t; tot += t
    for n, ct in sorted(agg.items(), key=lambda x: -x[1][1]):
        c, t = ct
        print(f"{n:70s} {c:5d} {t / 1e3:10.1f} us {100 * t / tot:5.1f}%  avg {t / c / 1e3:7.1f} us")
    print(f"total {tot / 1e3:.1f} us over {sum(c for c, _ in agg.values())} launches")


if __name__ == "__main__":
    main()
