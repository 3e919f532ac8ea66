#!/usr/bin/env python
"""DRAM traffic of the dominant kernel of a bench.py workload, as ncu sees it on bench.py's own batch 0.

  python scripts/ncu_traffic.py <tag> <kernel-regex> [bench.py flags ...]      # on the GPU box

1. `bench.py --one-forward <flags>` without ncu (must exit 0; prints the workload key, the experts batch 0 hit
   and the algorithmic bytes of that launch);
2. the same command under `ncu --set full --clock-control none --import-source on`, capturing the launch of
   <kernel-regex> that follows the warm-up forwards -> gpurun_out/full_<tag>.ncu-rep;
3. `ncu -i ... --page raw --csv` -> dram__bytes_read.sum + dram__bytes_write.sum, duration, tensor-pipe and
   DRAM utilisation of that launch, merged into gpurun_out/ncu_traffic_r2.json under the workload key
   (copied to profiles/ncu_traffic_r2.json, where bench.py's `roofline.traffic` reads it).
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = {"dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
        "gpu__time_duration.sum": "duration",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
        "sm__inst_executed_pipe_tensor.sum": "tensor_inst",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
        "lts__t_sector_hit_rate.pct": "l2_hit_pct", "lts__t_bytes.sum": "l2_bytes",
        "launch__registers_per_thread": "registers", "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3,
        "msecond": 1e3, "nsecond": 1e-3}


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, units, vals = rows[hdr], rows[hdr + 1], rows[hdr + 2]
    res = {"kernel_name": vals[names.index("Kernel Name")][:120]}
    for i, n in enumerate(names):
        if n in WANT:
            try:
                v = float(vals[i].replace(",", ""))
            except ValueError:
                continue
            res[WANT[n]] = v * UNIT.get(units[i], 1.0)
    return res


def main():
    tag, regex, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--one-forward", "--warmup", "3", *flags]
    plain = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if plain.returncode != 0:
        print("plain run failed:", plain.stderr[-800:])
        return 1
    info = json.loads([l for l in plain.stdout.splitlines() if l.startswith("{")][-1])
    rep = os.path.join(ROOT, "gpurun_out", f"full_{tag}.ncu-rep")
    ncu = ["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on", "--kernel-name-base", "demangled",
           "-k", f"regex:{regex}", "-s", str(info["warmup_forwards"]), "-c", "1", "-f", "-o", rep[:-8], *cmd]
    r = subprocess.run(ncu, capture_output=True, text=True, timeout=900)
    open(os.path.join(ROOT, "gpurun_out", f"ncu_full_{tag}.log"), "w").write(r.stdout[-4000:] + r.stderr[-2000:])
    if not os.path.exists(rep):
        print("no capture; ncu said:", (r.stdout + r.stderr)[-600:])
        return 1
    m = raw_metrics(rep)
    ent = {"kernel": info["kernel"], "kernel_name": m.get("kernel_name"), "experts_hit": info["experts_hit"],
           "algorithmic_bytes": info["algorithmic_bytes"], "dram_bytes": m.get("dram_read", 0) + m.get("dram_write", 0),
           "dram_read": m.get("dram_read"), "dram_write": m.get("dram_write"), "duration_us": m.get("duration"),
           "tensor_pipe_active_pct": m.get("tensor_pipe_active_pct"), "dram_throughput_pct": m.get("dram_throughput_pct"),
           "l2_hit_pct": m.get("l2_hit_pct"), "l2_bytes": m.get("l2_bytes"), "registers": m.get("registers"),
           "flops": info["flops"], "capture": f"full_{tag}.ncu-rep (ncu --set full --clock-control none; "
                                               f"bench.py --one-forward {' '.join(flags)})"}
    path = os.path.join(ROOT, "gpurun_out", "ncu_traffic_r2.json")
    allj = json.load(open(path)) if os.path.exists(path) else {}
    allj[info["one_forward"]] = ent
    json.dump(allj, open(path, "w"), indent=1)
    print(json.dumps({info["one_forward"]: ent}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
