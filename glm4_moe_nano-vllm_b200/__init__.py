"""B200-native GLM4-MoE sparse-expert block: drop-in for the reference's
`nanovllm/models/glm4_moe/moe.py` (`Glm4MoeMoE`, `GlmMoeSelectTopk`) backed by a C-ABI CUDA
library for sm_100a (`csrc/`, header `include/glm4_moe_b200.h`).  No CPU fallback."""
from . import synth  # noqa: F401

__all__ = ["synth", "Glm4MoeMoE", "GlmMoeSelectTopk", "Glm4MoeMLP", "Glm4MoeDenseMLP", "RMSNorm", "lib"]


def __getattr__(name):
    if name in ("Glm4MoeMoE", "GlmMoeSelectTopk", "Glm4MoeMLP", "Glm4MoeDenseMLP"):
        from . import moe
        return getattr(moe, name)
    if name == "RMSNorm":
        from . import layernorm
        return layernorm.RMSNorm
    if name == "lib":
        from . import _lib
        return _lib
    raise AttributeError(name)
