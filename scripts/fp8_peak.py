#!/usr/bin/env python
"""Dense e4m3 tensor-core peak of this box, measured the way MEASURED_PEAKS.json measures bf16:
cuBLASLt (torch._scaled_mm) e4m3 x e4m3 -> bf16, 8192^3, best of 10 (burst) and back to back for 4 s
(sustained).  Writes profiles/fp8_peak_r2.json, the denominator bench.py uses for the FP8 tier."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "fp8_peak_r2.json")
    dev = torch.device("cuda:0")
    N = 8192
    a = (torch.randn(N, N, device=dev) * 0.5).to(torch.float8_e4m3fn)
    b = (torch.randn(N, N, device=dev) * 0.5).to(torch.float8_e4m3fn).t()      # column-major second operand
    one = torch.ones((), device=dev)
    f = lambda: torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        f()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    flops = 2.0 * N ** 3
    n, t0 = 0, time.perf_counter()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    while time.perf_counter() - t0 < 4.0:
        for _ in range(20):
            f()
        n += 20
        torch.cuda.synchronize()
    e.record()
    torch.cuda.synchronize()
    sustained = flops * n / (s.elapsed_time(e) * 1e-3) / 1e12
    # bf16 on the same box for the ratio
    x = torch.randn(N, N, device=dev, dtype=torch.bfloat16)
    y = torch.randn(N, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        x @ y
    torch.cuda.synchronize()
    bb = 1e9
    for _ in range(10):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        x @ y
        e.record()
        torch.cuda.synchronize()
        bb = min(bb, s.elapsed_time(e))
    res = {"fp8_tflops": flops / (best * 1e-3) / 1e12, "fp8_tflops_sustained": sustained,
           "bf16_tflops_same_run": flops / (bb * 1e-3) / 1e12, "n": N,
           "how": "torch._scaled_mm e4m3 x e4m3 -> bf16 (cuBLASLt), 8192^3, best of 10 (burst) / back to back 4 s "
                  "(sustained), CUDA events", "gpu": torch.cuda.get_device_name(0), "torch": torch.__version__}
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
