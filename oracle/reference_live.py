"""The REAL reference module, imported from the installed copy.  TEST INFRASTRUCTURE ONLY.

`oracle/_ref/` holds the unmodified reference package (installed by `oracle/make_ref.py`); this
file imports `nanovllm.models.glm4_moe.moe` from it (or from `/root/reference` in the builder
container) so that

  * `tests/` can pin the restatement (`oracle/moe_oracle.py`) and the CUDA path against the
    reference itself on whatever machine they run on, GPU box included, and
  * `bench.py --impl reference` / `cpu_baseline` time the reference's own `Glm4MoeMoE.forward`
    (`moe.py:128-176`) on the box's host cores (`kind: "reference"`).

The single alteration: the reference's forward reads a debug safetensors file on
every call (`moe.py:153-157`, likewise `attention_new.py:288-292`) that exists only on its author's
machine; `safetensors.torch.load_file` is wrapped so that paths under `/data/ai_infra/debug` hand
back placeholders.  Every other path goes to the real `load_file`.
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = (os.path.join(ROOT, "oracle", "_ref"), "/root/reference")
_DEBUG_PREFIX = "/data/ai_infra/debug"


def reference_root() -> str | None:
    for c in CANDIDATES:
        if os.path.exists(os.path.join(c, "nanovllm", "models", "glm4_moe", "moe.py")):
            return c
    return None


def available() -> bool:
    return reference_root() is not None


def patch_debug_loads() -> None:
    """Neutralise the per-forward debug file reads (moe.py:153-157, attention_new.py:288-292)."""
    import safetensors.torch as st
    if getattr(st.load_file, "_glm4moe_patched", False):
        return
    real = st.load_file

    def load_file(filename, *a, **k):
        if str(filename).startswith(_DEBUG_PREFIX):
            z = torch.zeros(1)
            return {"experts_output": z, "shared_output": z, "final_hidden_states": z}
        return real(filename, *a, **k)

    load_file._glm4moe_patched = True
    st.load_file = load_file


def import_reference(root: str | None = None):
    """Put the reference package on sys.path (once) and return its `moe` module."""
    root = root or reference_root()
    if root is None:
        raise RuntimeError("the reference package is not installed: run `python oracle/make_ref.py` where "
                           "/root/reference is mounted (oracle/_ref/ then travels with the repo)")
    if root not in sys.path:
        sys.path.insert(0, root)
    patch_debug_loads()
    import nanovllm.models.glm4_moe.moe as ref_moe
    return ref_moe


def hf_config(shape):
    from transformers.models.glm4_moe import Glm4MoeConfig
    return Glm4MoeConfig(hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                         moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                         routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)


def build_reference_module(shape, dtype, device: str = "cpu"):
    """Instantiate the reference `Glm4MoeMoE` (moe.py:52-69) under the given default dtype / device,
    the way `ModelRunner` does (model_runner.py:29-31)."""
    ref_moe = import_reference()
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        with torch.device(device):
            moe = ref_moe.Glm4MoeMoE(hf_config(shape))
    finally:
        torch.set_default_dtype(prev)
    return moe


def build_reference_gate(shape, dtype, device: str = "cpu"):
    ref_moe = import_reference()
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        with torch.device(device):
            gate = ref_moe.GlmMoeSelectTopk(hf_config(shape))
    finally:
        torch.set_default_dtype(prev)
    return gate


def fill_reference_module(moe, weights) -> None:
    """Copy the stacked synthetic weights (synth.make_weights) into the reference's parameters."""
    E = len(moe.experts)
    with torch.no_grad():
        for e in range(E):
            moe.experts[e].gate_up_proj.weight.data.copy_(weights["w_gate_up"][e])
            moe.experts[e].down_proj.weight.data.copy_(weights["w_down"][e])
        moe.shared_experts.gate_up_proj.weight.data.copy_(weights["w_gate_up"][E])
        moe.shared_experts.down_proj.weight.data.copy_(weights["w_down"][E])
        moe.gate.weight.data.copy_(weights["gate_weight"])
        moe.gate.e_score_correction_bias.copy_(weights["bias"])
