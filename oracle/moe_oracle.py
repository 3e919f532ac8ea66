"""CPU ORACLE for the GLM4-MoE sparse-expert path.  TEST INFRASTRUCTURE ONLY.

This file is a torch-CPU restatement of the reference algorithm
(`/root/reference/nanovllm/models/glm4_moe/moe.py`, `mlp.py`, `layers/activation.py`,
`layers/linear.py`).  Only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py` may import it; the product path never does (and fails
loudly without its CUDA library).

Parity pin: `tests/golden/*.pt` were produced by importing the *real* reference module in the
builder container (`oracle/gen_golden.py`, which needs `/root/reference`) and this restatement
is checked bit-for-bit against them in `tests/test_oracle_golden.py`.  The reference ships no
golden vectors of its own (its fixtures live on the author's machine), so the
pin is "outputs of the reference itself, run here".

Third-party arithmetic: every op below executes inside torch (2.11.0 here; the reference
declares torch>=2.4, pyproject.toml:15).  `torch.topk(sorted=False)` order and tie-breaking
are implementation-defined; `tie_rule="lowest_index"` is the pinned rule of the CUDA router
(higher score first, then lower expert id) and equals `torch.topk` whenever the k-th and
(k+1)-th `scores_for_choice` differ, which the non-zero fp32 bias of §8d makes the normal case.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# router  (moe.py:28-50)
# --------------------------------------------------------------------------------------
def router_logits(h: torch.Tensor, gate_weight: torch.Tensor) -> torch.Tensor:
    """moe.py:32 — logits in the *model dtype* (not fp32)."""
    return F.linear(h.reshape(-1, h.shape[-1]), gate_weight)


def select_from_logits(logits: torch.Tensor, bias: torch.Tensor, k: int, scale: float,
                       tie_rule: str = "torch"):
    """moe.py:37-46 applied to given logits.  Returns (idx int64 [T,k], w model-dtype [T,k])."""
    scores = logits.sigmoid()                                        # moe.py:37 (model dtype)
    scores_for_choice = scores + bias.unsqueeze(0)                   # moe.py:38 (fp32: bias is fp32)
    if tie_rule == "torch":
        idx = torch.topk(scores_for_choice, k=k, dim=-1, sorted=False)[1]   # moe.py:39
    elif tie_rule == "lowest_index":
        idx = torch.sort(scores_for_choice, dim=-1, descending=True, stable=True)[1][:, :k]
    else:
        raise ValueError(tie_rule)
    w = scores.gather(1, idx)                                        # moe.py:41 (un-biased)
    denom = w.sum(dim=-1, keepdim=True) + 1e-20                      # moe.py:43
    w = w / denom                                                    # moe.py:44 (unconditional)
    w = w * scale                                                    # moe.py:46
    return idx, w


def router(h, gate_weight, bias, k, scale, tie_rule="torch"):
    return select_from_logits(router_logits(h, gate_weight), bias, k, scale, tie_rule)


def canonical_routing(idx: torch.Tensor, w: torch.Tensor):
    """Slot order inside a token is unspecified (sorted=False): sort slots by expert id."""
    order = torch.argsort(idx, dim=-1, stable=True)
    return idx.gather(1, order), w.gather(1, order)


# --------------------------------------------------------------------------------------
# implied permutation  (moe.py:135-144; SURVEY §8a-R5, probe A4)
# --------------------------------------------------------------------------------------
def permutation(idx: torch.Tensor, n_experts: int):
    """Expert-major, token-ascending order the reference loop visits (token, slot) pairs in.

    Returns (row_token int32 [T*k], row_slot int32 [T*k], expert_offsets int32 [E+1]).
    `row_token[r]` is the token feeding sorted row r; `expert_offsets[e]:expert_offsets[e+1]`
    are the rows of expert e.  Equivalent to a stable sort of the token-major flat list.
    """
    T, k = idx.shape
    flat = idx.reshape(-1).to(torch.int64)
    order = torch.sort(flat, stable=True)[1]
    counts = torch.bincount(flat, minlength=n_experts)
    offsets = torch.zeros(n_experts + 1, dtype=torch.int64)
    offsets[1:] = torch.cumsum(counts, 0)
    return (order // k).to(torch.int32), (order % k).to(torch.int32), offsets.to(torch.int32)


# --------------------------------------------------------------------------------------
# expert MLP  (mlp.py:49-54, activation.py:12-14, linear.py:74-75,151-152)
# --------------------------------------------------------------------------------------
def expert_mlp(x: torch.Tensor, w_gate_up: torch.Tensor, w_down: torch.Tensor) -> torch.Tensor:
    gate_up = F.linear(x, w_gate_up)            # [n, 2I]
    g, u = gate_up.chunk(2, -1)                 # first half = gate (activation.py:13)
    return F.linear(F.silu(g) * u, w_down)      # [n, H]


# --------------------------------------------------------------------------------------
# full block  (moe.py:128-176 minus the debug file read at :153-157)
# --------------------------------------------------------------------------------------
def moe_forward(h: torch.Tensor, weights: dict, k: int, scale: float, routing=None,
                tie_rule: str = "torch", compute_dtype=None) -> dict:
    """weights: stacked dict from synth.make_weights (index E of the stacks = shared expert).

    `routing=(idx, w)` forces the routing (used for the fp32 truth run).
    `compute_dtype` upcasts tokens and each expert's weights on the fly (fp32 truth without
    materialising an fp32 copy of a multi-GB stack).
    """
    w_gu, w_d = weights["w_gate_up"], weights["w_down"]
    E = w_gu.shape[0] - 1
    h2 = h.reshape(-1, h.shape[-1])                                  # moe.py:130
    if routing is None:
        idx, w = router(h2, weights["gate_weight"], weights["bias"], k, scale, tie_rule)
    else:
        idx, w = routing
    cast = (lambda t: t) if compute_dtype is None else (lambda t: t.to(compute_dtype))
    h2 = cast(h2)
    w = cast(w)
    acc = torch.zeros_like(h2)                                       # moe.py:133
    for e in range(E):                                               # moe.py:135
        mask = idx == e
        if not mask.any():
            continue
        tok, pos = torch.where(mask)                                 # moe.py:142-143
        y = expert_mlp(h2[tok], cast(w_gu[e]), cast(w_d[e]))         # moe.py:147-148
        acc.index_add_(0, tok, y * w[tok, pos].unsqueeze(-1))        # moe.py:149-151
    shared = expert_mlp(h2, cast(w_gu[E]), cast(w_d[E]))             # moe.py:165
    return {"out": acc + shared, "experts_out": acc, "shared_out": shared,   # moe.py:171
            "idx": idx, "w": w}


def moe_forward_fp32(h: torch.Tensor, weights: dict, k: int, scale: float, routing) -> dict:
    """fp32 truth: model-dtype values upcast to fp32, routing forced to the native one."""
    return moe_forward(h, weights, k, scale, routing=routing, compute_dtype=torch.float32)


# --------------------------------------------------------------------------------------
# quant tier (BASELINE.json configs[2]).  The reference has no FP8 path; this restates the
# published DeepSeek-V3-style block format the tier implements on top of the reference's
# structure (mlp.py:49-54): weights e4m3 with one fp32 scale per 128x128 block (dequantised by
# the caller), activations e4m3 with one fp32 scale per token per 128 columns, scale = amax/448,
# products exact, fp32 accumulation; the SiLU*mul intermediate is rounded to the model dtype
# (it is stored in it) before it is re-quantised for the down projection.
# `pow2_scales` (UE8M0 variant of the same published format, the scale type the Blackwell
# block-scaled MMA reads): every activation scale is rounded UP to a power of two; weight block
# scales are powers of two already when they reach the oracle (dequantised by the caller).
# --------------------------------------------------------------------------------------
def fp8_act_roundtrip(a: torch.Tensor, pow2_scales: bool = False) -> torch.Tensor:
    """a [n, K] -> fp32 values of its per-token 1x128 e4m3 quantisation."""
    n, K = a.shape
    g = a.float().reshape(n, K // 128, 128)
    scale = g.abs().amax(-1, keepdim=True).clamp_min(1e-10) / 448.0
    if pow2_scales:
        m, e = torch.frexp(scale)                       # scale = m * 2^e, m in [0.5, 1)
        scale = torch.ldexp(torch.ones_like(m), torch.where(m == 0.5, e - 1, e))
    q = (g / scale).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).float()
    return (q * scale).reshape(n, K)


def expert_mlp_fp8(x: torch.Tensor, w_gate_up_deq: torch.Tensor, w_down_deq: torch.Tensor,
                   model_dtype: torch.dtype, pow2_scales: bool = False) -> torch.Tensor:
    gate_up = F.linear(fp8_act_roundtrip(x, pow2_scales), w_gate_up_deq.float())
    g, u = gate_up.chunk(2, -1)
    act = (F.silu(g) * u).to(model_dtype)
    return F.linear(fp8_act_roundtrip(act, pow2_scales), w_down_deq.float())


def moe_forward_fp8(h: torch.Tensor, deq_weights: dict, routing, model_dtype: torch.dtype,
                    pow2_scales: bool = False) -> torch.Tensor:
    """fp32 evaluation of the quant-tier algorithm with forced routing (idx, w)."""
    w_gu, w_d = deq_weights["w_gate_up"], deq_weights["w_down"]
    E = w_gu.shape[0] - 1
    h2 = h.reshape(-1, h.shape[-1])
    idx, w = routing
    acc = torch.zeros(h2.shape, dtype=torch.float32)
    for e in range(E):
        mask = idx == e
        if not mask.any():
            continue
        tok, pos = torch.where(mask)
        y = expert_mlp_fp8(h2[tok], w_gu[e], w_d[e], model_dtype, pow2_scales)
        acc.index_add_(0, tok, y * w[tok, pos].float().unsqueeze(-1))
    return acc + expert_mlp_fp8(h2, w_gu[E], w_d[E], model_dtype, pow2_scales)


def count_experts_hit(idx: torch.Tensor) -> int:
    return int(torch.unique(idx).numel())


# --------------------------------------------------------------------------------------
# the step in front of the block: residual add + RMSNorm
# --------------------------------------------------------------------------------------
def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """layers/layernorm.py:17-26 (`rms_forward`) and :29-40 (`add_rms_forward`), out of place."""
    orig = x.dtype
    x32 = x.float()
    res_out = None
    if residual is not None:
        x32 = x32 + residual.float()                        # layernorm.py:35
        res_out = x32.to(orig)                              # :36
    var = x32.pow(2).mean(dim=-1, keepdim=True)             # :22 / :37
    y = (x32 * torch.rsqrt(var + eps)).to(orig) * weight    # :23-24 / :38-39
    return y if residual is None else (y, res_out)


# --------------------------------------------------------------------------------------
# the tensor-parallel boundary in front of the block
# --------------------------------------------------------------------------------------
def allreduce_add_rmsnorm(partials, weight, eps: float, residual: torch.Tensor | None = None):
    """What decode_layer.py:96-98 computes from the per-rank o_proj outputs: `dist.all_reduce(y)` of
    `RowParallelLinear.forward` (layers/linear.py:153-154: the sum lands in the model dtype), then
    `post_attention_layernorm(y, residual)` (layers/layernorm.py:29-40).  The sum order of an all-reduce is the
    library's business; pinned here: fp32, ascending rank, one rounding (for two ranks that IS what NCCL computes:
    a + b is commutative and is rounded once).  `weight=None`: the all-reduce alone."""
    acc = torch.zeros_like(partials[0], dtype=torch.float32)
    for p in partials:
        acc = acc + p.float()
    y = acc.to(partials[0].dtype)
    if weight is None:
        return y
    return rmsnorm(y, weight, eps, residual)

