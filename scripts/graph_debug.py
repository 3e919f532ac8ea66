"""eager vs eager vs graph replay on one module (debug aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config
shape, dt = synth.MINI, torch.bfloat16
wts = synth.make_weights(shape, dt)
torch.set_default_dtype(dt)
moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape))
torch.set_default_dtype(torch.float32)
moe.load_stacked(wts)
moe = moe.cuda()
xa = synth.make_tokens(shape, 16, dt, 4).cuda()
xb = synth.make_tokens(shape, 16, dt, 5).cuda()
e1 = moe(xa); torch.cuda.synchronize()
e2 = moe(xa); torch.cuda.synchronize()
e3 = moe(xb); torch.cuda.synchronize()
e4 = moe(xa); torch.cuda.synchronize()
print("eager a1==a2", torch.equal(e1, e2), "a1==a4", torch.equal(e1, e4), "a==b", torch.equal(e1, e3))
static_in = xa.clone()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    so = moe(static_in)
for name, src, ref in (("b", xb, e3), ("a", xa, e1), ("b", xb, e3), ("a", xa, e1)):
    static_in.copy_(src); g.replay(); torch.cuda.synchronize()
    print("replay", name, torch.equal(so, ref), float((so.float() - ref.float()).abs().max()))
    ee = moe(static_in); torch.cuda.synchronize()
    print("  eager after", torch.equal(ee, ref))
