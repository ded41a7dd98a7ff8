"""top source lines by stall samples + a few raw metrics from ncu csv exports.  usage: ncu_top.py src.csv raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[2])))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ("gpu__time_duration.sum", "lts__t_requests_srcunit_tex_op_read.sum", "smsp__inst_executed.sum", "dram__bytes_write.sum", "dram__bytes_read.sum",
        "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "launch__grid_size")
for h, u, v in zip(hdr, units, vals):
    if h in want: print(h, u, v)
rows = list(csv.reader(open(sys.argv[1])))
agg = []; fname = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split('/')[-1]; continue
    if len(r) < 8 or r[0] == "Line No": continue
    if r[0] != "":
        try: agg.append((int(r[4]), fname, r[0], r[1].strip()[:100], r[7]))
        except ValueError: pass
agg.sort(reverse=True)
print("total samples", sum(a[0] for a in agg))
for a in agg[:int(sys.argv[3]) if len(sys.argv) > 3 else 14]: print(a)
