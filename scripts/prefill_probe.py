"""A few prefill (T=4096) forwards of the Air layer; used under ncu for the grouped-GEMM profile."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config
T = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
shape, dt = synth.AIR, torch.bfloat16
torch.set_default_dtype(dt)
moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=T)
torch.set_default_dtype(torch.float32)
moe = moe.cuda()
torch.manual_seed(0)
with torch.no_grad():
    moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
    moe.gate.e_score_correction_bias.uniform_(0, 0.1)
xs = [torch.randn(T, shape.H, device="cuda", dtype=dt) for _ in range(2)]
for i in range(n):
    y = moe(xs[i % 2])
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
