"""CPU (needs the reference: /root/reference or its installed copy oracle/_ref/): the drop-in boundary of SURVEY 8b exercised
against the REAL reference classes.  The reference decoder layer (decode_layer.py:10-38) is built once
with its own `Glm4MoeMoE` and once with ours bound to the same name (`from .moe import Glm4MoeMoE`,
decode_layer.py:6), and the two sparse blocks must be indistinguishable to their caller: same
state_dict keys / shapes / dtypes, and `load_weights(path, prefix)` (decode_layer.py:75 ->
moe.py:71-126, mlp.py:83-131) filling identical values from one HF-layout checkpoint."""
import os
import sys
import types

import pytest
import torch

import glm4_moe_b200
from glm4_moe_b200 import synth

from oracle import reference_live as R

pytestmark = pytest.mark.skipif(not R.available(), reason="reference package neither mounted nor installed")


def _config(shape):
    # a plain attribute bag: the reference only reads attributes (transformers 5.5 renamed rope_theta /
    # rope_scaling, which attention_new.py:90-95 still expects)
    return types.SimpleNamespace(
        hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k, moe_intermediate_size=shape.I,
        intermediate_size=4 * shape.H, n_shared_experts=1, n_group=1, topk_group=1,
        routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True, num_attention_heads=8,
        num_key_value_heads=2, head_dim=64, first_k_dense_replace=1, num_hidden_layers=2,
        max_position_embeddings=512, rope_theta=10000.0, rope_scaling=None, partial_rotary_factor=0.5,
        rms_norm_eps=1e-5, attention_bias=False, hidden_act="silu", use_qk_norm=False)


def _layers(shape, dtype):
    R.import_reference()
    from nanovllm.models.glm4_moe import decode_layer as DL
    from nanovllm.models.glm4_moe.moe import Glm4MoeMoE as RefMoE
    cfg = _config(shape)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        DL.Glm4MoeMoE = RefMoE
        ref = DL.Glm4MoeDecoderLayer(cfg, "model.layers.1")
        DL.Glm4MoeMoE = glm4_moe_b200.Glm4MoeMoE          # the one-line swap of INTEGRATION.md
        ours = DL.Glm4MoeDecoderLayer(cfg, "model.layers.1")
    finally:
        DL.Glm4MoeMoE = RefMoE
        torch.set_default_dtype(prev)
    return ref, ours


def test_swap_inside_reference_decoder_layer():
    shape = synth.MINI
    ref, ours = _layers(shape, torch.bfloat16)
    assert type(ref.mlp).__module__.startswith("nanovllm.")
    assert isinstance(ours.mlp, glm4_moe_b200.Glm4MoeMoE)
    a, b = ref.mlp.state_dict(), ours.mlp.state_dict()
    assert set(a) == set(b)
    for k, v in a.items():
        assert tuple(b[k].shape) == tuple(v.shape) and b[k].dtype == v.dtype, k
    assert b["gate.e_score_correction_bias"].dtype == torch.float32           # moe.py:26
    # dense layers (layer_idx < first_k_dense_replace) are not touched by the swap (decode_layer.py:31-36)
    assert [n for n, _ in ref.named_children()] == [n for n, _ in ours.named_children()]


def test_load_weights_matches_reference_loader(tmp_path):
    from safetensors.torch import save_file
    shape, dt = synth.MoeShape(256, 8, 2, 128, 1.0, "loader-parity"), torch.bfloat16
    w = synth.make_weights(shape, dt)
    p = "model.layers.1.mlp"
    sd = {f"{p}.gate.weight": w["gate_weight"], f"{p}.gate.e_score_correction_bias": w["bias"],
          "model.layers.1.input_layernorm.weight": torch.ones(shape.H, dtype=dt),
          # a sibling prefix that merely STARTS with ours must not leak in (the reference's startswith
          # match, mlp.py:96, is replaced by an exact `prefix + "."` match)
          "model.layers.11.mlp.gate.weight": torch.zeros_like(w["gate_weight"])}
    for e in range(shape.E + 1):
        q = f"{p}.experts.{e}" if e < shape.E else f"{p}.shared_experts"
        sd[f"{q}.gate_proj.weight"] = w["w_gate_up"][e, :shape.I].contiguous()
        sd[f"{q}.up_proj.weight"] = w["w_gate_up"][e, shape.I:].contiguous()
        sd[f"{q}.down_proj.weight"] = w["w_down"][e].contiguous()
    os.makedirs(tmp_path / "ckpt")
    save_file(sd, str(tmp_path / "ckpt" / "model-00001-of-00001.safetensors"))
    ref, ours = _layers(shape, dt)
    ref.mlp.load_weights(str(tmp_path / "ckpt"), p)          # the reference's own loader
    ours.mlp.load_weights(str(tmp_path / "ckpt"), p)
    a, b = ref.mlp.state_dict(), ours.mlp.state_dict()
    for k, v in a.items():
        assert torch.equal(b[k], v), k
    with pytest.raises(ValueError):                            # moe.py:94-95
        ours.mlp.load_weights(str(tmp_path / "ckpt"), "model.layers.7.mlp")
