"""Router token-tile rows at prefill (GLM4MOE_KNOB_ROUTER_ROWS): router stage time and whole-layer time
for each setting, both shapes, one process.  Usage: router_probe.py [T ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import glm4_moe_b200
from glm4_moe_b200 import synth, _lib
from bench import _hf_config
Ts = [int(a) for a in sys.argv[1:]] or [4096, 8192]
dt = torch.bfloat16
lib = _lib.load()
for sname in ("glm-4.5-air", "glm-4.5"):
    shape = synth.SHAPES[sname]
    torch.manual_seed(0)
    torch.set_default_dtype(dt)
    moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=max(Ts))
    torch.set_default_dtype(torch.float32)
    moe = moe.cuda()
    with torch.no_grad():
        moe._w_gate_up.normal_(0, 0.02); moe._w_down.normal_(0, 0.02); moe.gate.weight.normal_(0, 0.02)
        moe.gate.e_score_correction_bias.uniform_(0, 0.1)
    for T in Ts:
        xs = [torch.randn(T, shape.H, device="cuda", dtype=dt) for _ in range(4)]
        ref_idx = None
        for rows in (0, 32, 64, 128):
            _lib.check(lib.glm4moe_debug_set(3, rows))
            for i in range(3):
                moe(xs[i % 4])
            acc, n = 0.0, 16
            for i in range(n):
                _, ms = moe.forward_timed(xs[i % 4])
                acc += ms["router_logits"] / n
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); s0.record()
            for i in range(n):
                moe(xs[i % 4])
            s1.record(); torch.cuda.synchronize()
            moe(xs[0])
            idx = moe.workspace_view("idx", T).clone()
            lg = moe.workspace_view("logits", T).clone()
            if ref_idx is None:
                ref_idx, ref_lg = idx, lg
            print(json.dumps({"shape": sname, "T": T, "router_rows": rows, "router_us": round(acc * 1e3, 1),
                              "layer_ms": round(s0.elapsed_time(s1) / n, 4),
                              "logits_bit_identical": bool(torch.equal(lg, ref_lg)),
                              "idx_identical": bool(torch.equal(idx, ref_idx))}), flush=True)
    _lib.check(lib.glm4moe_debug_set(3, 0))
    del moe
    torch.cuda.empty_cache()
