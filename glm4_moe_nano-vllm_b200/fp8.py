"""Offline FP8 conversion of expert weights: e4m3 values + fp32 scales per 128x128 block
(BASELINE.json configs[2]; the DeepSeek-V3-style block format, `weight_scale_inv` in HF checkpoints).

`w ~= q.float() * scale[block]`, `scale = amax(block) / 448`.  Host-side preparation only (torch
ops); the hot path consumes the result through `glm4moe_layer_create(..., s_gate_up, s_down, ...)`.
"""
from __future__ import annotations

import torch

E4M3_MAX = 448.0
BLOCK = 128


def pow2_ceil(s: torch.Tensor) -> torch.Tensor:
    """Smallest power of two >= s (fp32, s > 0): the UE8M0 rounding of a scale."""
    m, e = torch.frexp(s.float())                       # s = m * 2^e, m in [0.5, 1)
    return torch.ldexp(torch.ones_like(m), torch.where(m == 0.5, e - 1, e))


def is_pow2(s: torch.Tensor) -> bool:
    """Every scale a normal power of two (what GLM4MOE_W_FP8_BLOCK128_POW2 requires)."""
    s = s.float()
    return bool(((torch.frexp(s)[0] == 0.5) & (s >= 2.0 ** -126) & (s <= 2.0 ** 127)).all())


def quantize_block128(w: torch.Tensor, pow2_scales: bool = False):
    """w [..., R, C] (R, C multiples of 128) -> (q float8_e4m3fn [..., R, C], scale fp32 [..., R/128, C/128]).
    `pow2_scales`: scales rounded up to a power of two (UE8M0 checkpoints; the hardware block-scaled MMA)."""
    *lead, R, C = w.shape
    if R % BLOCK or C % BLOCK:
        raise ValueError(f"block-128 FP8 needs dimensions divisible by 128, got {R}x{C}")
    blocks = w.float().reshape(*lead, R // BLOCK, BLOCK, C // BLOCK, BLOCK)
    amax = blocks.abs().amax(dim=(-3, -1), keepdim=True).clamp_min(1e-10)
    scale = amax / E4M3_MAX
    if pow2_scales:
        scale = pow2_ceil(scale)
    q = (blocks / scale).clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn)
    return q.reshape(*lead, R, C), scale.reshape(*lead, R // BLOCK, C // BLOCK).contiguous()


def dequantize_block128(q: torch.Tensor, scale: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    *lead, R, C = q.shape
    blocks = q.float().reshape(*lead, R // BLOCK, BLOCK, C // BLOCK, BLOCK)
    return (blocks * scale.reshape(*lead, R // BLOCK, 1, C // BLOCK, 1)).reshape(*lead, R, C).to(dtype)


def quantize_rows_1x128(x: torch.Tensor, pow2_scales: bool = False):
    """Reference for the device activation quantiser: per-token groups of 128 along the last dim."""
    *lead, K = x.shape
    g = x.float().reshape(*lead, K // BLOCK, BLOCK)
    scale = g.abs().amax(-1, keepdim=True).clamp_min(1e-10) / E4M3_MAX
    if pow2_scales:
        scale = pow2_ceil(scale)
    q = (g / scale).clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn)
    return q.reshape(*lead, K), scale.squeeze(-1)


def convert_checkpoint(src: str, dst: str, prefixes=None, pow2_scales: bool = False) -> dict:
    """Offline conversion of the sparse-block expert weights of an HF-layout checkpoint to the block
    FP8 format `Glm4MoeMoE.load_weights` reads.

    `src`: a `.safetensors` file or a directory of them; `dst`: output file / directory (same file
    names).  Every `...experts.{i}.{gate,up,down}_proj.weight` and `...shared_experts.*_proj.weight`
    (optionally restricted to keys starting with one of `prefixes`) is replaced by an e4m3 `weight`
    and an fp32 `weight_scale_inv [rows/128, cols/128]`; router weights, the fp32
    `e_score_correction_bias` and everything outside the sparse block are copied through untouched.
    One pass over the keys; one output FILE is held in memory at a time (safetensors writes a file whole).  Returns {"converted": n, "copied": m}."""
    import os
    import re
    from safetensors import safe_open
    from safetensors.torch import save_file
    pat = re.compile(r"\.(experts\.\d+|shared_experts)\.(gate_proj|up_proj|down_proj)\.weight$")
    if os.path.isdir(src):
        os.makedirs(dst, exist_ok=True)
        pairs = [(os.path.join(src, f), os.path.join(dst, f)) for f in sorted(os.listdir(src))
                 if f.endswith(".safetensors")]
    else:
        pairs = [(src, dst)]
    stats = {"converted": 0, "copied": 0}
    for fin, fout in pairs:
        out = {}
        with safe_open(fin, framework="pt", device="cpu") as f:
            for key in f.keys():
                t = f.get_tensor(key)
                wanted = prefixes is None or any(key.startswith(p + ".") for p in prefixes)
                if wanted and pat.search(key) and t.dtype != torch.float8_e4m3fn:
                    q, s = quantize_block128(t.contiguous(), pow2_scales)
                    out[key] = q
                    out[key[:-len("weight")] + "weight_scale_inv"] = s
                    stats["converted"] += 1
                else:
                    out[key] = t.contiguous()
                    stats["copied"] += 1
        save_file(out, fout)
    return stats


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description="bf16/fp16 GLM4-MoE checkpoint -> block-128 FP8 expert weights")
    ap.add_argument("src")
    ap.add_argument("dst")
    ap.add_argument("--prefix", action="append", help="only convert keys under this prefix (repeatable)")
    ap.add_argument("--pow2-scales", action="store_true", help="power-of-two (UE8M0) block scales")
    a = ap.parse_args()
    print(convert_checkpoint(a.src, a.dst, a.prefix, a.pow2_scales))
