"""Pivot an ncu --csv metrics log (one row per kernel x metric) into a table. Usage: kern_table.py file.csv"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        hdr, start = r, i
        break
ki, mi, vi, ii = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
ui = hdr.index("Metric Unit")
per = collections.OrderedDict()
for r in rows[start + 1:]:
    if len(r) <= vi:
        continue
    d = per.setdefault(r[ii], {"kernel": r[ki]})
    d[r[mi]] = (r[vi], r[ui])
short = lambda n: n.split("(")[0].replace("glm4moe::", "").replace("void ", "")[:60]
for id_, d in per.items():
    g = lambda m: d.get(m, ("", ""))
    print(f"{id_:>4} {short(d['kernel']):60s} t={g('gpu__time_duration.sum')[0]:>10}{g('gpu__time_duration.sum')[1]:3s} "
          f"rd={g('dram__bytes_read.sum')[0]:>9}{g('dram__bytes_read.sum')[1]:6s} wr={g('dram__bytes_write.sum')[0]:>9}{g('dram__bytes_write.sum')[1]:6s} "
          f"dram%={g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')[0]:>6} tc%={g('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active')[0]:>6} "
          f"l2hit={g('lts__t_sector_hit_rate.pct')[0]:>6} grid={g('launch__grid_size')[0]}")
