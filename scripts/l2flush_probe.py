"""Decode latency with the L2 flushed between iterations (SURVEY 8d timing protocol, T <= 4): a 512 MB
buffer is overwritten before every forward, each forward is timed alone with its own event pair."""
import json, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config
shape, dt = synth.AIR, torch.bfloat16
fp8 = "--fp8" in sys.argv
torch.manual_seed(0)
torch.set_default_dtype(dt)
moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=256)
torch.set_default_dtype(torch.float32)
moe = moe.cuda()
with torch.no_grad():
    moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
    moe.gate.e_score_correction_bias.uniform_(0, 0.1)
if fp8:
    moe.quantize_fp8(drop_native=True)
scratch = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for T in (1, 2, 4, 32):
    xs = [torch.randn(T, shape.H, device="cuda", dtype=dt) for _ in range(16)]
    with torch.inference_mode():
        for i in range(4):
            moe(xs[i])
        res = {}
        for flush in (False, True):
            ts = []
            for i in range(40):
                if flush:
                    scratch.fill_(i & 0xff)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); moe(xs[i % 16]); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3)
            ts.sort()
            res["flushed" if flush else "rotating"] = {"median_us": round(statistics.median(ts), 1),
                                                       "p10_us": round(ts[4], 1), "p90_us": round(ts[35], 1)}
    print(json.dumps({"fp8": fp8, "T": T, **res}), flush=True)
