"""Per-stage timings (CUDA events between stages) of the Air layer for a list of T."""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config
FP8 = "--fp8" in sys.argv
Ts = [int(a) for a in sys.argv[1:] if not a.startswith("--")] or [1, 4, 32, 256, 4096]
shape, dt = synth.SHAPES[os.environ.get("GLM4MOE_SHAPE", "glm-4.5-air")], torch.bfloat16
torch.manual_seed(0)
torch.set_default_dtype(dt)
moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=max(Ts))
torch.set_default_dtype(torch.float32)
moe = moe.cuda()
with torch.no_grad():
    moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
    moe.gate.e_score_correction_bias.uniform_(0, 0.1)
if FP8:
    moe.quantize_fp8(drop_native=True)
for T in Ts:
    xs = [torch.randn(T, shape.H, device="cuda", dtype=dt) for _ in range(8)]
    for i in range(5):
        moe(xs[i % 8])
    acc = {}
    n = 30
    for i in range(n):
        _, ms = moe.forward_timed(xs[i % 8])
        for k, v in ms.items():
            acc[k] = acc.get(k, 0) + v / n
    # plain (no events between stages)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s0.record()
    for i in range(n):
        moe(xs[i % 8])
    s1.record(); torch.cuda.synchronize()
    # graph replay
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(8):
            moe(xs[i])
    g.replay(); torch.cuda.synchronize()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for i in range(4):
        g.replay()
    g1.record(); torch.cuda.synchronize()
    print(json.dumps({"fp8": FP8, "T": T, "eager_ms": s0.elapsed_time(s1) / n, "graph_ms": g0.elapsed_time(g1) / 32,
                      "stages_us": {k: round(v * 1e3, 1) for k, v in acc.items()}}))
