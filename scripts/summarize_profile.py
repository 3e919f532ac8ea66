#!/usr/bin/env python
"""Turn gpurun_out/launches_<tag>.csv + gpurun_out/prof_<tag>.ncu-rep into profiles/<tag>.md."""
import collections
import csv
import io
import subprocess
import sys

# usage: summarize_profile.py <tag> [full-capture .ncu-rep ...]   (default capture: gpurun_out/prof_<tag>.ncu-rep)
tag = sys.argv[1]
reps = sys.argv[2:] or [f"gpurun_out/prof_{tag}.ncu-rep"]
out = [f"# ncu summary `{tag}`", ""]
try:
    lines = [l for l in open(f"gpurun_out/launches_{tag}.csv") if not l.startswith("==")]
    agg, n = collections.OrderedDict(), 0
    for row in csv.DictReader(io.StringIO("".join(lines))):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
        a = agg.setdefault(row["Kernel Name"][:90], [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(a[1] for a in agg.values())
    out += [f"## launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`): {n} launches, "
            f"{tot:.0f} us total (cold-cache, serialised: compare shares)", "",
            "| total us | launches | avg us | share | kernel |", "|---:|---:|---:|---:|---|"]
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        out.append(f"| {a[1]:.1f} | {a[0]} | {a[1] / a[0]:.2f} | {100 * a[1] / tot:.1f}% | `{k}` |")
    out.append("")
except FileNotFoundError:
    pass
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg"]
for rep in reps:
    try:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        r = list(csv.reader(io.StringIO(raw)))
        hdr = r[0]
        ki = hdr.index("Kernel Name")
        out += [f"## full capture `{rep}` (`ncu --set full --clock-control none --import-source on`)", ""]
        names = [row[ki][:80] for row in r[2:]]
        out.append("| metric | unit | " + " | ".join(f"`{n}`" for n in names) + " |")
        out.append("|---|---|" + "---|" * len(names))
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                out.append(f"| {w} | {r[1][i]} | " + " | ".join(row[i] for row in r[2:]) + " |")
        out.append("")
    except Exception as e:  # noqa: BLE001
        out.append(f"(no full capture {rep}: {e})")
open(f"profiles/{tag}.md", "w").write("\n".join(out) + "\n")
print("\n".join(out))
