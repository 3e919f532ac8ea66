"""Offline FP8 conversion of expert weights: e4m3 values + fp32 scales per 128x128 block
(BASELINE.json configs[2]; the DeepSeek-V3-style block format, `weight_scale_inv` in HF checkpoints).

`w ~= q.float() * scale[block]`, `scale = amax(block) / 448`.  Host-side preparation only (torch
ops); the hot path consumes the result through `glm4moe_layer_create(..., s_gate_up, s_down, ...)`.
"""
from __future__ import annotations

import torch

E4M3_MAX = 448.0
BLOCK = 128


def quantize_block128(w: torch.Tensor):
    """w [..., R, C] (R, C multiples of 128) -> (q float8_e4m3fn [..., R, C], scale fp32 [..., R/128, C/128])."""
    *lead, R, C = w.shape
    if R % BLOCK or C % BLOCK:
        raise ValueError(f"block-128 FP8 needs dimensions divisible by 128, got {R}x{C}")
    blocks = w.float().reshape(*lead, R // BLOCK, BLOCK, C // BLOCK, BLOCK)
    amax = blocks.abs().amax(dim=(-3, -1), keepdim=True).clamp_min(1e-10)
    scale = amax / E4M3_MAX
    q = (blocks / scale).clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn)
    return q.reshape(*lead, R, C), scale.reshape(*lead, R // BLOCK, C // BLOCK).contiguous()


def dequantize_block128(q: torch.Tensor, scale: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    *lead, R, C = q.shape
    blocks = q.float().reshape(*lead, R // BLOCK, BLOCK, C // BLOCK, BLOCK)
    return (blocks * scale.reshape(*lead, R // BLOCK, 1, C // BLOCK, 1)).reshape(*lead, R, C).to(dtype)


def quantize_rows_1x128(x: torch.Tensor):
    """Reference for the device activation quantiser: per-token groups of 128 along the last dim."""
    *lead, K = x.shape
    g = x.float().reshape(*lead, K // BLOCK, BLOCK)
    scale = g.abs().amax(-1, keepdim=True).clamp_min(1e-10) / E4M3_MAX
    q = (g / scale).clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn)
    return q.reshape(*lead, K), scale.squeeze(-1)
