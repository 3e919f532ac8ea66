#!/usr/bin/env python
"""Dense layer-0 MLP (GLM-4.5-Air: H=4096, I=10944) through `glm4moe_mlp_forward` against the reference's own
layers on the same GPU (`F.linear` -> SiLU*mul -> `F.linear`, cuBLAS; mlp.py:49-54).  One JSON line per T."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from glm4_moe_b200.moe import Glm4MoeDenseMLP

H, I, dt = 4096, 10944, torch.bfloat16
torch.set_default_dtype(dt)
m = Glm4MoeDenseMLP(H, I, "silu", tp_rank=0, tp_size=1)
torch.set_default_dtype(torch.float32)
m = m.cuda()
with torch.no_grad():
    m.gate_up_proj.weight.normal_(0, 0.02)
    m.down_proj.weight.normal_(0, 0.02)
wgu, wd = m.gate_up_proj.weight, m.down_proj.weight


def ref(x):
    g, u = F.linear(x, wgu).chunk(2, -1)
    return F.linear(F.silu(g) * u, wd)


def timed(f, x, n):
    for _ in range(5):
        f(x)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        f(x)
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


for T in [int(a) for a in sys.argv[1:]] or [1, 32, 256, 4096]:
    x = torch.randn(T, H, device="cuda", dtype=dt)
    with torch.inference_mode():
        y, yr = m(x), ref(x)
        rel = float((y.float() - yr.float()).norm() / yr.float().norm())
        n = 200 if T <= 256 else 30
        ours, cublas = timed(m, x, n), timed(ref, x, n)
    by = 3 * H * I * 2 + T * (2 * H + 2 * I) * 2
    fl = 6.0 * T * H * I
    print(json.dumps({"what": "dense MLP H=4096 I=10944 bf16", "T": T, "ours_ms": ours, "reference_layers_cublas_ms": cublas,
                      "ours_GBps": by / ours / 1e6, "ours_TFLOPs": fl / ours / 1e9, "rel_diff_vs_cublas_path": rel}))
