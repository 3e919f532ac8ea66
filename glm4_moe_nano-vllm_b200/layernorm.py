"""Drop-in for the reference's `nanovllm/layers/layernorm.py` RMSNorm (the step in front of the sparse
block: `post_attention_layernorm`, `models/glm4_moe/decode_layer.py:98`), backed by
`glm4moe_add_rmsnorm` (csrc/rmsnorm.cu).  Same constructor, parameter name and call forms:
`norm(x) -> y` and `norm(x, residual) -> (y, residual_out)`; no CPU fallback."""
from __future__ import annotations

import torch
from torch import nn

from . import _lib


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """Functional form of `RMSNorm.forward` (`glm4moe_add_rmsnorm`): `y`, or `(y, residual_out)` with a residual."""
    _lib.require_cuda(x, "x")
    H = x.shape[-1]
    if x.dtype != weight.dtype:
        raise _lib.Glm4MoeLibraryError(f"x dtype {x.dtype} != norm weight dtype {weight.dtype}")
    xc = x.contiguous()
    T = xc.numel() // H
    out = torch.empty_like(xc)
    res_out = None
    rp = None
    if residual is not None:
        rc = residual.contiguous()
        if rc.shape != xc.shape or rc.dtype != xc.dtype:
            raise _lib.Glm4MoeLibraryError("residual must match x in shape and dtype")
        res_out = torch.empty_like(xc)
        rp = rc.data_ptr()
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().glm4moe_add_rmsnorm(xc.data_ptr(), rp, weight.data_ptr(), float(eps), T, H,
                                                  _lib.DTYPE_CODE[x.dtype], out.data_ptr(), _lib.ptr(res_out),
                                                  _lib.stream_ptr(x.device)), "add_rmsnorm")
    return out if residual is None else (out, res_out)


class RMSNorm(nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-6) -> None:     # layernorm.py:7-14
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size))

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None):
        """layernorm.py:42-50.  Unlike the reference's in-place `add_`/`mul_` on temporaries, inputs are
        left untouched and fresh tensors are returned."""
        return rmsnorm(x, self.weight, self.eps, residual)
