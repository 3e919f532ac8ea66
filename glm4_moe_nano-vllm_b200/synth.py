"""Deterministic synthetic shards for the GLM4-MoE sparse-expert path.

Everything is drawn on the CPU with a seeded ``torch.Generator`` in fp32 and then cast, so the
same bytes are produced in the builder container, on the GPU box's host and in the golden
generator.  Draw order follows the reference module's ``named_parameters()`` order
(``nanovllm/models/glm4_moe/moe.py:56-69``): experts.0.gate_up_proj, experts.0.down_proj, ...,
gate.weight, shared_experts.gate_up_proj, shared_experts.down_proj, then the fp32 bias buffer.
"""
from __future__ import annotations

import dataclasses
import hashlib

import torch

WEIGHT_SEED = 1234
TOKEN_SEED = 4321
WEIGHT_STD = 0.02
BIAS_HIGH = 0.1


@dataclasses.dataclass(frozen=True)
class MoeShape:
    """Shape of one sparse-expert block (names follow transformers' Glm4MoeConfig)."""

    hidden_size: int
    n_routed_experts: int
    num_experts_per_tok: int
    moe_intermediate_size: int
    routed_scaling_factor: float = 1.0
    name: str = "custom"

    @property
    def H(self) -> int:
        return self.hidden_size

    @property
    def E(self) -> int:
        return self.n_routed_experts

    @property
    def k(self) -> int:
        return self.num_experts_per_tok

    @property
    def I(self) -> int:  # noqa: E743
        return self.moe_intermediate_size

    def flops_per_token(self) -> int:
        """Algorithmic flops per token: router + k routed + 1 shared."""
        H, E, k, I = self.H, self.E, self.k, self.I
        return 2 * H * E + k * 6 * H * I + 6 * H * I

    def algorithmic_bytes(self, T: int, experts_hit: int, weight_bytes: int = 2, act_bytes: int = 2,
                          fp8_scales: bool = False) -> int:
        """Algorithmic HBM bytes of one forward over T tokens."""
        H, E, I = self.H, self.E, self.I
        per_expert = 3 * H * I * weight_bytes
        if fp8_scales:
            per_expert += 4 * (-(-2 * I // 128) * -(-H // 128) + -(-H // 128) * -(-I // 128))
        return (experts_hit + 1) * per_expert + E * H * act_bytes + 2 * T * H * act_bytes


TINY = MoeShape(1024, 16, 4, 512, 1.0, "tiny")            # BASELINE.json configs[0]
AIR = MoeShape(4096, 128, 8, 1408, 1.0, "glm-4.5-air")     # configs[1..3]
GLM45 = MoeShape(5120, 160, 8, 1536, 2.5, "glm-4.5")       # configs[4]
# reduced shapes for fast parity cases (same divisibility properties as the named ones)
MINI = MoeShape(512, 32, 8, 256, 2.5, "mini")
SMALL = MoeShape(2048, 32, 8, 704, 1.0, "small")           # SURVEY probe A6 shape (I = 5.5 * 128)

SHAPES = {s.name: s for s in (TINY, AIR, GLM45, MINI, SMALL)}


def _randn(gen: torch.Generator, shape, dtype: torch.dtype, std: float) -> torch.Tensor:
    # chunked so the fp32 staging copy of a 4.4 GB stack never exists at once
    out = torch.empty(shape, dtype=dtype)
    flat = out.view(-1)
    step = 1 << 24
    for s in range(0, flat.numel(), step):
        n = min(step, flat.numel() - s)
        flat[s:s + n] = (torch.randn(n, generator=gen, dtype=torch.float32) * std).to(dtype)
    return out


def make_weights(shape: MoeShape, dtype: torch.dtype, seed: int = WEIGHT_SEED,
                 zero_bias: bool = False) -> dict[str, torch.Tensor]:
    """Stacked synthetic weights.

    Returns ``w_gate_up [E+1, 2I, H]`` and ``w_down [E+1, H, I]`` (index E = shared expert),
    ``gate_weight [E, H]`` in ``dtype`` and ``bias [E]`` fp32.  Row ``[0, I)`` of ``w_gate_up[e]``
    is gate_proj and ``[I, 2I)`` is up_proj (``mlp.py:118``).
    """
    H, E, I = shape.H, shape.E, shape.I
    gen = torch.Generator(device="cpu").manual_seed(seed)
    w_gu = torch.empty((E + 1, 2 * I, H), dtype=dtype)
    w_d = torch.empty((E + 1, H, I), dtype=dtype)
    for e in range(E):
        w_gu[e] = _randn(gen, (2 * I, H), dtype, WEIGHT_STD)
        w_d[e] = _randn(gen, (H, I), dtype, WEIGHT_STD)
    gate_w = _randn(gen, (E, H), dtype, WEIGHT_STD)
    w_gu[E] = _randn(gen, (2 * I, H), dtype, WEIGHT_STD)
    w_d[E] = _randn(gen, (H, I), dtype, WEIGHT_STD)
    bias = torch.rand(E, generator=gen, dtype=torch.float32) * BIAS_HIGH
    if zero_bias:
        bias.zero_()
    return {"w_gate_up": w_gu, "w_down": w_d, "gate_weight": gate_w, "bias": bias}


def make_weights_device(shape: MoeShape, dtype: torch.dtype, device, experts=None, seed: int = WEIGHT_SEED):
    """Same distributions as `make_weights`, drawn on the GPU with one generator seed per expert so
    an expert-parallel rank can create just its shard (bench.py; values differ from `make_weights`).
    Returns stacks over `experts` (default all) followed by the shared expert."""
    H, E, I = shape.H, shape.E, shape.I
    experts = list(range(E)) if experts is None else list(experts)
    n = len(experts)
    w_gu = torch.empty((n + 1, 2 * I, H), dtype=dtype, device=device)
    w_d = torch.empty((n + 1, H, I), dtype=dtype, device=device)
    gen = torch.Generator(device=device)
    for slot, e in enumerate(experts + [E]):
        gen.manual_seed(seed * 1000 + e)
        w_gu[slot] = (torch.randn((2 * I, H), generator=gen, device=device, dtype=torch.float32) * WEIGHT_STD).to(dtype)
        w_d[slot] = (torch.randn((H, I), generator=gen, device=device, dtype=torch.float32) * WEIGHT_STD).to(dtype)
    gen.manual_seed(seed * 1000 + E + 1)
    gate_w = (torch.randn((E, H), generator=gen, device=device, dtype=torch.float32) * WEIGHT_STD).to(dtype)
    bias = torch.rand(E, generator=gen, device=device, dtype=torch.float32) * BIAS_HIGH
    return {"w_gate_up": w_gu, "w_down": w_d, "gate_weight": gate_w, "bias": bias}


def make_router_weights(shape: MoeShape, dtype: torch.dtype, seed: int = WEIGHT_SEED + 77):
    """Router-only shard (gate weight [E,H] + fp32 bias) for the big router shapes."""
    gen = torch.Generator(device="cpu").manual_seed(seed)
    gate_w = (torch.randn((shape.E, shape.H), generator=gen, dtype=torch.float32) * WEIGHT_STD).to(dtype)
    bias = torch.rand(shape.E, generator=gen, dtype=torch.float32) * BIAS_HIGH
    return gate_w, bias


def make_tokens(shape: MoeShape, T: int, dtype: torch.dtype, batch_index: int = 0,
                seed: int = TOKEN_SEED) -> torch.Tensor:
    gen = torch.Generator(device="cpu").manual_seed(seed + batch_index)
    return torch.randn((T, shape.H), generator=gen, dtype=torch.float32).to(dtype)


def tensor_sha256(t: torch.Tensor) -> str:
    """Checksum of the raw bytes, used by golden fixtures to pin regenerated inputs."""
    t = t.detach().contiguous().cpu()
    return hashlib.sha256(t.view(torch.uint8).numpy().tobytes()).hexdigest()


def expected_experts_hit(shape: MoeShape, T: int) -> float:
    E, k = shape.E, shape.k
    return E * (1.0 - (1.0 - k / E) ** T)
