"""Stage-by-stage triage of the FP8 tier at a small shape: act rows and y rows vs a torch fp32
restatement of the quant-tier algorithm, row by row (debug aid, not a test)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import glm4_moe_b200
from glm4_moe_b200 import synth
from bench import _hf_config
from oracle import moe_oracle as O

shape = synth.SHAPES[os.environ.get("GLM4MOE_SHAPE", "mini")]
dt = torch.bfloat16
Ts = [int(a) for a in sys.argv[1:]] or [1, 40]
wts = synth.make_weights(shape, dt)
torch.set_default_dtype(dt)
moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=256)
torch.set_default_dtype(torch.float32)
moe.load_stacked(wts)
moe = moe.cuda()
moe.quantize_fp8()
deq = moe.dequantized_weights()
wgu, wd = deq["w_gate_up"].cpu().float(), deq["w_down"].cpu().float()
E, k, I, H = shape.E, shape.k, shape.I, shape.H


def rt_kernel(a):
    """the device quantiser's arithmetic: scale = amax * (1/448), q = v * (1/scale)"""
    n, K = a.shape
    g = a.float().reshape(n, K // 128, 128)
    scale = g.abs().amax(-1, keepdim=True).clamp_min(1e-10) * torch.tensor(1.0 / 448.0, dtype=torch.float32)
    inv = 1.0 / scale
    q = (g * inv).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).float()
    return (q * scale).reshape(n, K)


for T in Ts:
    x = synth.make_tokens(shape, T, dt, T)
    out = moe(x.cuda())
    torch.cuda.synchronize()
    offs = moe.workspace_view("expert_offsets").cpu()
    row_token = moe.workspace_view("row_token").cpu()
    act = moe.workspace_view("act").cpu().float()
    y = moe.workspace_view("y").cpu().float()
    rw = moe.workspace_view("row_weight").cpu()
    tiles = moe.workspace_view("tiles").cpu()
    nt = int(moe.workspace_view("num_tiles")[0])
    print(f"== T={T} tiles={nt} tile_rows={tiles[:nt].tolist()[:12]}")
    worst = []
    tot = {"a": [0, 0], "a2": [0, 0], "y": [0, 0], "y2": [0, 0], "x12": [0, 0]}
    for e in list(range(E)) + [E]:
        if e < E:
            r0, r1 = int(offs[e]), int(offs[e + 1])
            toks = row_token[r0:r1].long()
            rows = list(range(r0, r1))
            wts_r = rw[r0:r1]
        else:
            toks = torch.arange(T)
            rows = list(range(T * k, T * k + T))
            wts_r = torch.ones(T)
        if not rows:
            continue
        xe = x[toks]
        gu = F.linear(O.fp8_act_roundtrip(xe), wgu[e])
        g, u = gu.chunk(2, -1)
        a_ref = (F.silu(g) * u).to(dt).float()
        a_got = act[rows]
        ea = (a_got - a_ref).abs().max().item() / (a_ref.abs().max().item() + 1e-9)
        # down from the kernel's own act (isolates the second GEMM)
        y_ref = F.linear(O.fp8_act_roundtrip(a_got.to(dt)), wd[e]) * wts_r[:, None]
        y_got = y[rows]
        ey = (y_got - y_ref).abs().max().item() / (y_ref.abs().max().item() + 1e-9)
        gu2 = F.linear(rt_kernel(xe), wgu[e])
        g2, u2 = gu2.chunk(2, -1)
        a_ref2 = (F.silu(g2) * u2).to(dt).float()
        y_ref2 = F.linear(rt_kernel(a_got.to(dt)), wd[e]) * wts_r[:, None]
        for key, got_, ref_ in (("a", a_got, a_ref), ("a2", a_got, a_ref2), ("y", y_got, y_ref), ("y2", y_got, y_ref2),
                                ("x12", O.fp8_act_roundtrip(xe), rt_kernel(xe))):
            tot[key][0] += (got_ - ref_).pow(2).sum().item()
            tot[key][1] += ref_.pow(2).sum().item()
        worst.append((max(ea, ey), e, len(rows), rows[0], ea, ey))
        if ea > 2e-2 or ey > 2e-2:
            # which rows / which feature ranges are off
            da = ((a_got - a_ref).abs().amax(1) / (a_ref.abs().amax(1) + 1e-9))
            dy = ((y_got - y_ref).abs().amax(1) / (y_ref.abs().amax(1) + 1e-9))
            fa = (a_got - a_ref).abs().reshape(len(rows), -1, 128).amax(-1).amax(0) / (a_ref.abs().max() + 1e-9)
            fy = (y_got - y_ref).abs().reshape(len(rows), -1, 128).amax(-1).amax(0) / (y_ref.abs().max() + 1e-9)
            print(f"expert {e} rows {rows[0]}..{rows[-1]} ea={ea:.3g} ey={ey:.3g}")
            print("   act err per row :", [round(v, 3) for v in da.tolist()])
            print("   y   err per row :", [round(v, 3) for v in dy.tolist()])
            print("   act err per ftile:", [round(v, 3) for v in fa.tolist()])
            print("   y   err per ftile:", [round(v, 3) for v in fy.tolist()])
    print("rel-F per stage:", {k_: round((v[0] / max(v[1], 1e-30)) ** 0.5, 5) for k_, v in tot.items()})
    worst.sort(reverse=True)
    print("worst:", [(round(a, 4), e, n, r0, round(b, 4), round(c, 4)) for a, e, n, r0, b, c in worst[:5]])
    idx = moe.workspace_view("idx", T).cpu().long()
    w = moe.workspace_view("w", T).cpu().to(dt)
    algo = O.moe_forward_fp8(x, {"w_gate_up": wgu, "w_down": wd}, (idx, w), dt)
    d = out.float().cpu() - algo
    print("final rel-F", (d.norm() / algo.norm()).item(), "max", d.abs().max().item(), "rms", algo.pow(2).mean().sqrt().item())
