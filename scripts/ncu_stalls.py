"""Warp-stall samples of a warp-specialised kernel split by the mbarrier each spin loop waits on.

All mbar_wait() call sites inline to the same source line, so the ncu source page cannot tell the
TMA producer's wait (empty slot) from the MMA issuer's (operands / accumulator) — but the SASS
`SYNCS.PHASECHK.TRANS64.TRYWAIT P, [R + URZ + 0xOFFSET]` carries the barrier's shared-memory offset.
Usage: python scripts/ncu_stalls.py report.ncu-rep [--stages 3 --stage-bytes 65536 --warps 20]
(barrier block layout of umma_gemm.cu: full[stages], empty[stages], tmem_full[2], tmem_empty[2],
sched_full[4], sched_empty[4], 8 bytes each, right after the operand ring)."""
import argparse
import csv
import io
import re
import subprocess

ap = argparse.ArgumentParser()
ap.add_argument("report")
ap.add_argument("--stages", type=int, default=3)
ap.add_argument("--stage-bytes", type=int, default=65536)
ap.add_argument("--warps", type=int, default=20, help="warps per CTA (to express samples as a share of one warp's time)")
a = ap.parse_args()
raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
h = rows[hdr]
isrc, isamp = h.index("Source"), h.index("# Samples")
data = [(int(r[isamp] or 0), r[isrc]) for r in rows[hdr + 1:] if len(r) > isamp]
tot = sum(s for s, _ in data)
base = a.stages * a.stage_bytes
names = {}
for i in range(a.stages):
    names[base + 8 * i] = f"full[{i}]: MMA warp waits for operands"
    names[base + 8 * (a.stages + i)] = f"empty[{i}]: TMA producer waits for a free slot"
o = base + 16 * a.stages
for i in range(2):
    names[o + 8 * i] = f"tmem_full[{i}]: epilogue warps wait for the accumulator"
    names[o + 16 + 8 * i] = f"tmem_empty[{i}]: MMA warp waits for the epilogue"
for i in range(4):
    names[o + 32 + 8 * i] = f"sched_full[{i}]"
    names[o + 64 + 8 * i] = f"sched_empty[{i}]"
agg = {}
for i, (s, src) in enumerate(data):
    m = re.search(r"TRYWAIT P\d, \[\w+\+URZ(?:\+(0x[0-9a-f]+))?\]", src)
    if m:
        off = int(m.group(1) or "0x0", 16)
        agg[off] = agg.get(off, 0) + sum(d[0] for d in data[max(0, i - 1):i + 2])   # YIELD + TRYWAIT + BRA
per_warp = tot / a.warps
print(f"{tot} samples, {a.warps} warps per CTA")
for off, v in sorted(agg.items()):
    print(f"  mbarrier +{off:#x} {names.get(off, '(register-indexed: first of its group)')}: {v} samples = "
          f"{100 * v / tot:.1f} % of all = {100 * v / per_warp:.0f} % of one warp's time")
mma = sum(s for s, src in data if "UTCHMMA" in src or "UTCBAR" in src)
print(f"  UTCHMMA / UTCBAR issue (tensor-pipe back-pressure): {mma} samples = {100 * mma / per_warp:.0f} % of one warp's time")
