"""A few forwards of the Air layer at one T (ncu target: `ncu -k regex:glm4moe ... python scripts/one_forward.py 1`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config
FP8 = "--fp8" in sys.argv or "--fp8-pow2" in sys.argv
POW2 = "--fp8-pow2" in sys.argv
args = [int(a) for a in sys.argv[1:] if not a.startswith("--")]
T = args[0] if args else 32
iters = args[1] if len(args) > 1 else 6
shape = synth.SHAPES[os.environ.get("GLM4MOE_SHAPE", "glm-4.5-air")]
dt = torch.bfloat16
torch.manual_seed(0)
torch.set_default_dtype(dt)
moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=T)
torch.set_default_dtype(torch.float32)
moe = moe.cuda()
with torch.no_grad():
    moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
    moe.gate.e_score_correction_bias.uniform_(0, 0.1)
if FP8:
    moe.quantize_fp8(drop_native=True, pow2_scales=POW2)
xs = [torch.randn(T, shape.H, device="cuda", dtype=dt) for _ in range(iters)]
torch.cuda.synchronize()
for x in xs:
    moe(x)
torch.cuda.synchronize()
print("done", T, moe.last_impl, moe.last_launches)
