"""Decode through a STACK of GLM-4.5-Air sparse blocks (BASELINE.json configs[3] without attention):
L x [residual add + RMSNorm -> MoE block], every layer with its own weights, replayed from one CUDA
graph.  Unlike the single-layer bench no weight can sit in L2 between steps and every layer boundary
(PDL chain, graph replay) is inside the timed region.  Usage: stack_probe.py [--fp8] [--layers L] T...
45 FP8 layers = 101 GB (one B200); bf16 holds 30 layers (134 GB)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config

import argparse
ap = argparse.ArgumentParser()
ap.add_argument("--fp8", action="store_true")
ap.add_argument("--layers", type=int, default=0)
ap.add_argument("tokens", type=int, nargs="*", default=[1, 32])
A = ap.parse_args()
FP8, Ts = A.fp8, A.tokens
L = A.layers or (45 if FP8 else 30)
shape, dt = synth.AIR, torch.bfloat16
torch.manual_seed(0)
layers, norms = [], []
for i in range(L):
    torch.set_default_dtype(dt)
    moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=max(Ts))
    norm = glm4_moe_b200.RMSNorm(shape.H)
    torch.set_default_dtype(torch.float32)
    moe, norm = moe.cuda(), norm.cuda()
    with torch.no_grad():
        moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
        moe.gate.e_score_correction_bias.uniform_(0, 0.1)
    if FP8:
        moe.quantize_fp8(drop_native=True)
    layers.append(moe); norms.append(norm)
torch.cuda.synchronize()
bytes_w = (3 * shape.H * shape.I) * (1 if FP8 else 2)
for T in Ts:
    x = torch.randn(T, shape.H, device="cuda", dtype=dt)

    def step(h):
        res = None
        for moe, norm in zip(layers, norms):
            if res is None:
                hn, res = norm(h), h
            else:
                hn, res = norm(h, res)
            h = moe(hn)
        return h, res

    with torch.inference_mode():
        for _ in range(2):
            step(x)
        torch.cuda.synchronize()
        hit = [int(torch.unique(m.workspace_view("idx", T)).numel()) for m in layers]
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out, _ = step(x)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n):
            g.replay()
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    U = sum(hit) / len(hit)
    by = L * ((U + 1) * bytes_w + shape.E * shape.H * 2 + 2 * T * shape.H * 2)
    print(json.dumps({"fp8": FP8, "layers": L, "T": T, "ms_per_step": ms, "us_per_layer": ms * 1e3 / L,
                      "tokens_per_s": T / ms * 1e3, "experts_hit_mean": U, "GBps": by / ms / 1e6,
                      "finite": bool(torch.isfinite(out.float()).all())}))
