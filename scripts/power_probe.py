#!/usr/bin/env python
"""Board power and SM clock while a workload loops for a few seconds (NVML averages power over ~0.1 s, so a
bench step is far too short to see it).  Answers one question: are the prefill GEMMs at the power cap?

Workloads: cuBLAS bf16 8192^3 (what MEASURED_PEAKS.json's tensor peak is), a pure HBM copy, our dense MLP at
T=4096 (compute only), the Air sparse block at T=4096 (compute + 4.5 GB of weights per step), at T=32 (HBM only).
One JSON line each: ms per call, achieved TF/s and GB/s, median / max watts, the enforced limit, median SM MHz."""
import json
import os
import statistics
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import glm4_moe_b200
from glm4_moe_b200 import synth
from glm4_moe_b200.moe import Glm4MoeDenseMLP

import pynvml

pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
LIMIT = pynvml.nvmlDeviceGetEnforcedPowerLimit(h) / 1000.0


class Sampler:
    def __init__(self):
        self.w, self.mhz, self.stop = [], [], threading.Event()
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop.is_set():
            self.w.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
            self.mhz.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
            time.sleep(0.02)


def loop(name, fn, flops, bytes_, seconds=3.0):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        fn()
    e.record()
    torch.cuda.synchronize()
    n = max(20, int(seconds * 1000 / (s.elapsed_time(e) / 10)))
    sm = Sampler()
    time.sleep(1.0)                       # idle gap: the previous workload's heat / averaging window
    sm.t.start()
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    sm.stop.set()
    sm.t.join()
    ms = s.elapsed_time(e) / n
    half = len(sm.w) // 2                 # second half of the window: the averaging filter has settled
    print(json.dumps({"workload": name, "ms": round(ms, 4), "TFLOPs": round(flops / ms / 1e9, 1),
                      "GBps": round(bytes_ / ms / 1e6, 1), "watts_median": round(statistics.median(sm.w[half:]), 1),
                      "watts_max": round(max(sm.w), 1), "limit_w": LIMIT,
                      "sm_mhz_median": statistics.median(sm.mhz[half:]), "calls": n}), flush=True)


dt = torch.bfloat16
with torch.inference_mode():
    a = torch.randn(8192, 8192, device="cuda", dtype=dt)
    b = torch.randn(8192, 8192, device="cuda", dtype=dt)
    loop("cuBLAS bf16 8192^3", lambda: torch.mm(a, b), 2 * 8192.0 ** 3, 3 * 8192 * 8192 * 2)
    src = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    dst = torch.empty_like(src)
    loop("device copy 1 GiB", lambda: dst.copy_(src), 0, 2.0 * (1 << 30))
    del a, b, src, dst

    H, I = 4096, 10944
    torch.set_default_dtype(dt)
    m = Glm4MoeDenseMLP(H, I, "silu", tp_rank=0, tp_size=1)
    torch.set_default_dtype(torch.float32)
    m = m.cuda()
    m.gate_up_proj.weight.normal_(0, 0.02)
    m.down_proj.weight.normal_(0, 0.02)
    x = torch.randn(4096, H, device="cuda", dtype=dt)
    loop("dense MLP T=4096 (ours)", lambda: m(x), 6.0 * 4096 * H * I, 3 * H * I * 2)
    del m

    sys.argv = [sys.argv[0]]
    import bench
    shape = synth.SHAPES["glm-4.5-air"]
    for weights in ("bf16", "fp8_pow2"):
        torch.set_default_dtype(dt)
        torch.set_default_device("cuda")
        moe = glm4_moe_b200.Glm4MoeMoE(bench._hf_config(shape), max_tokens=4096)
        torch.set_default_device("cpu")
        torch.set_default_dtype(torch.float32)
        moe.load_local_stacks(synth.make_weights_device(shape, dt, torch.device("cuda"), experts=range(0, shape.E)))
        if weights != "bf16":
            moe.quantize_fp8(drop_native=True, pow2_scales=True)
        elt = 2 if weights == "bf16" else 1
        for T in (4096, 32):
            xs = [synth.make_tokens(shape, T, dt, 50 + i).cuda() for i in range(4)]
            it = [0]

            def step():
                it[0] += 1
                return moe(xs[it[0] % 4])
            hit = shape.E if T >= 1024 else 102
            loop(f"Air sparse block {weights} T={T}", step, 2.0 * T * (shape.k + 1) * 3 * shape.H * shape.I,
                 (hit + 1) * 3 * shape.H * shape.I * elt)
        del moe
        torch.cuda.empty_cache()
