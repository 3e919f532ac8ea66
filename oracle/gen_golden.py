"""Generate golden fixtures by running the REAL reference module (builder container only).

Needs the reference (`/root/reference`, or its installed copy `oracle/_ref/`, see
`oracle/make_ref.py` and `oracle/reference_live.py`).  Usage:  python oracle/gen_golden.py
Writes `tests/golden/<case>.pt` holding: the case description (shape, dtype, T, seeds), sha256
checksums of the regenerated inputs, and the reference's outputs (`gate()` indices + weights,
`experts_output`, `shared_output`, final output) — the same three tensors the reference's own
debug dump names (moe.py:160-172).

The reference's `forward` reads a debug safetensors file on every call (moe.py:153-157) that
exists only on its author's machine; `safetensors.torch.load_file` is wrapped to hand back
placeholders for those paths.  Nothing else of the reference is altered.
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def _load_pkg():
    sys.path.insert(0, ROOT)
    import glm4_moe_b200  # noqa: F401  (import shim at the repo root)
    from glm4_moe_b200 import synth
    return synth


sys.path.insert(0, ROOT)
from oracle.reference_live import (build_reference_gate, build_reference_module,   # noqa: E402,F401
                                   fill_reference_module, import_reference)


CASES = [
    # name, shape name, dtype, T, batch_index, zero_bias
    ("tiny_fp32_T64", "tiny", torch.float32, 64, 0, False),        # BASELINE.json configs[0]
    ("mini_bf16_T48", "mini", torch.bfloat16, 48, 0, False),
    ("mini_fp16_T33", "mini", torch.float16, 33, 1, False),
    ("mini_bf16_T1", "mini", torch.bfloat16, 1, 2, False),
    ("small_bf16_T256", "small", torch.bfloat16, 256, 0, False),   # SURVEY probe A6 shape
]
ROUTER_ONLY = [
    # Air / GLM-4.5 router shape only (full weights would be GBs): gate() outputs
    ("air_router_bf16_T256", "glm-4.5-air", torch.bfloat16, 256, 0),
    ("glm45_router_bf16_T64", "glm-4.5", torch.bfloat16, 64, 0),
]


def main():
    synth = _load_pkg()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    torch.set_num_threads(8)
    for name, sname, dtype, T, bi, zero_bias in CASES:
        shape = synth.SHAPES[sname]
        weights = synth.make_weights(shape, dtype, zero_bias=zero_bias)
        x = synth.make_tokens(shape, T, dtype, bi)
        moe = build_reference_module(shape, dtype)
        fill_reference_module(moe, weights)
        with torch.inference_mode():
            idx, w = moe.gate(x)
            final = moe(x)                      # moe.py:128-176
            shared = moe.shared_experts(x)      # the "shared_output" of moe.py:165-166
        blob = {
            "case": {"name": name, "shape": sname, "dtype": str(dtype), "T": T, "batch_index": bi,
                     "zero_bias": zero_bias, "weight_seed": synth.WEIGHT_SEED,
                     "token_seed": synth.TOKEN_SEED, "torch": torch.__version__},
            "sha": {n: synth.tensor_sha256(t) for n, t in weights.items()} | {"x": synth.tensor_sha256(x)},
            "idx": idx.to(torch.int32), "w": w, "shared_out": shared, "out": final,
        }
        torch.save(blob, os.path.join(out_dir, name + ".pt"))
        print(name, "out", tuple(final.shape), final.dtype, "rms", float(final.float().pow(2).mean().sqrt()))
    for name, sname, dtype, T, bi in ROUTER_ONLY:
        shape = synth.SHAPES[sname]
        gate_w, bias = synth.make_router_weights(shape, dtype)
        x = synth.make_tokens(shape, T, dtype, bi)
        gate = build_reference_gate(shape, dtype)
        with torch.no_grad():
            gate.weight.data.copy_(gate_w)
            gate.e_score_correction_bias.copy_(bias)
        with torch.inference_mode():
            idx, w = gate(x)
        blob = {"case": {"name": name, "shape": sname, "dtype": str(dtype), "T": T, "batch_index": bi,
                         "gate_seed": synth.WEIGHT_SEED + 77, "torch": torch.__version__},
                "sha": {"gate_weight": synth.tensor_sha256(gate_w), "bias": synth.tensor_sha256(bias),
                        "x": synth.tensor_sha256(x)},
                "idx": idx.to(torch.int32), "w": w}
        torch.save(blob, os.path.join(out_dir, name + ".pt"))
        print(name, tuple(idx.shape))
    # RMSNorm in front of the block (layers/layernorm.py), with and without the residual
    import_reference()
    from nanovllm.layers.layernorm import RMSNorm
    for name, H, T, dtype in (("rmsnorm_bf16_H4096_T5", 4096, 5, torch.bfloat16), ("rmsnorm_fp32_H1024_T3", 1024, 3, torch.float32),
                              ("rmsnorm_fp16_H512_T9", 512, 9, torch.float16)):
        g = torch.Generator().manual_seed(99)
        x = (torch.randn(T, H, generator=g) * 1.5).to(dtype)
        res = torch.randn(T, H, generator=g).to(dtype)
        wgt = (1.0 + 0.1 * torch.randn(H, generator=g)).to(dtype)
        prev = torch.get_default_dtype()
        torch.set_default_dtype(dtype)
        try:
            norm = RMSNorm(H, eps=1e-5)
        finally:
            torch.set_default_dtype(prev)
        with torch.no_grad():
            norm.weight.copy_(wgt)
            y0 = norm(x.clone())
            y1, r1 = norm(x.clone(), res.clone())
        torch.save({"case": {"name": name, "H": H, "T": T, "dtype": str(dtype), "eps": 1e-5, "seed": 99},
                    "x": x, "residual": res, "weight": wgt, "y_plain": y0, "y_add": y1, "residual_out": r1},
                   os.path.join(out_dir, name + ".pt"))
        print(name, float(y1.float().abs().mean()))


if __name__ == "__main__":
    main()
