"""Routing skew (SURVEY 8d "skew set", 8e "per-rank load imbalance under uniform and Zipf routing").

Part 1 (CPU, always): expert popularity and per-rank row imbalance (max rows on a rank / mean) of the
contiguous expert blocks of distributed.py at R = 2/4/8, for the uniform synthetic bias and for a
Zipf(1.2)-shaped per-expert offset added to e_score_correction_bias, using the oracle router on the
router-only weights of both shapes.
Part 2 (GPU, if present): layer time of the Air shape at T = 32 / 256 / 4096 under both biases.
Usage: python scripts/skew_probe.py [--gpu]"""
import json, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth
from oracle import moe_oracle as O      # checker / analysis only (not on the product path)

ZIPF_S, STRENGTH = 1.2, 0.5


def zipf_offset(E: int) -> torch.Tensor:
    """Offset in [-STRENGTH, 0]: expert of popularity rank r gets STRENGTH * (-s ln(r+1)) / (s ln E); ranks are a
    fixed pseudo-random permutation of the expert ids so hot experts are spread over the rank blocks."""
    g = torch.Generator().manual_seed(7)
    rank = torch.randperm(E, generator=g).float()
    return STRENGTH * (-(ZIPF_S * torch.log(rank + 1.0)) / (ZIPF_S * math.log(E)))


def imbalance(idx: torch.Tensor, E: int):
    cnt = torch.bincount(idx.reshape(-1), minlength=E).float()
    out = {"rows_per_expert_max_over_mean": round(float(cnt.max() / cnt.mean()), 2),
           "experts_hit": int((cnt > 0).sum())}
    for R in (2, 4, 8):
        per = cnt.reshape(R, E // R).sum(1)
        out[f"rank_max_over_mean_R{R}"] = round(float(per.max() / per.mean()), 3)
    return out


def main():
    dt = torch.bfloat16
    for sname, Ts in (("glm-4.5-air", (32, 256, 4096)), ("glm-4.5", (64, 8192))):
        shape = synth.SHAPES[sname]
        gate_w, bias = synth.make_router_weights(shape, dt)
        for T in Ts:
            x = synth.make_tokens(shape, T, dt, 0)
            for name, b in (("uniform", bias), ("zipf1.2", bias + zipf_offset(shape.E))):
                idx, _ = O.router(x, gate_w, b, shape.k, shape.routed_scaling_factor, tie_rule="lowest_index")
                print(json.dumps({"part": "imbalance", "shape": sname, "T": T, "bias": name, **imbalance(idx.long(), shape.E)}),
                      flush=True)
    if "--gpu" not in sys.argv or not torch.cuda.is_available():
        return
    from bench import _hf_config
    shape = synth.AIR
    torch.manual_seed(0)
    torch.set_default_dtype(dt)
    moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=4096)
    torch.set_default_dtype(torch.float32)
    moe = moe.cuda()
    with torch.no_grad():
        moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
        base = torch.rand(shape.E, device="cuda") * 0.1
    for name, b in (("uniform", base), ("zipf1.2", base + zipf_offset(shape.E).cuda())):
        with torch.no_grad():
            moe.gate.e_score_correction_bias.copy_(b)
        for T in (32, 256, 4096):
            xs = [torch.randn(T, shape.H, device="cuda", dtype=dt) for _ in range(8)]
            with torch.inference_mode():
                for i in range(4):
                    moe(xs[i])
                n = 50 if T <= 256 else 12
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize(); s0.record()
                for i in range(n):
                    moe(xs[i % 8])
                s1.record(); torch.cuda.synchronize()
                idx = moe.workspace_view("idx", T).cpu().long()
            print(json.dumps({"part": "gpu", "shape": shape.name, "T": T, "bias": name, "ms": round(s0.elapsed_time(s1) / n, 4),
                              **imbalance(idx, shape.E)}), flush=True)


if __name__ == "__main__":
    main()
