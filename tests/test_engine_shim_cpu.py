"""CPU: the construction-time shim that lets the unmodified reference engine host this package's sparse
block (`glm4_moe_b200.engine_shim`).  The forward itself needs flash-attn and the CUDA library
(tests/test_gpu_engine.py); here: the synthetic model directory satisfies the reference's `Config` and
`AutoTokenizer`, the attribute view restores what transformers 5.x renamed, all layers are built with the
swap in place, the residual reaches the final norm, and the dummy-weight path fills every parameter."""
import os

import pytest
import torch
import torch.distributed as dist

import glm4_moe_b200
from glm4_moe_b200 import engine_shim as S

pytestmark = pytest.mark.skipif(S.reference_root() is None, reason="reference engine not installed (oracle/make_ref.py)")

SMALL = dict(hidden_size=256, n_routed_experts=8, num_experts_per_tok=2, moe_intermediate_size=128,
             intermediate_size=512, num_hidden_layers=3, num_attention_heads=4, num_key_value_heads=2,
             head_dim=64, vocab_size=512, routed_scaling_factor=2.5)


@pytest.fixture(scope="module")
def group():
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
        dist.init_process_group("gloo", rank=0, world_size=1)
    yield
    dist.destroy_process_group()


def test_model_dir_config_and_tokenizer(tmp_path):
    S.install()
    d = S.make_model_dir(str(tmp_path / "m"), **SMALL)
    from nanovllm.config import Config
    cfg = Config(d, max_num_batched_tokens=512, max_model_len=256)
    hf = cfg.hf_config
    assert isinstance(hf, S.HfConfigView)
    assert hf.rope_theta == 10000.0 and hf.rope_scaling is None and hf.head_dim == 64
    assert hf.torch_dtype == torch.bfloat16 and hf.model_type == "glm4_moe"
    assert hf.num_hidden_layers == 3 and hf.n_routed_experts == 8 and hf.first_k_dense_replace == 1
    from transformers import AutoTokenizer
    tok = AutoTokenizer.from_pretrained(d, use_fast=True)
    assert tok.eos_token_id == 1
    assert tok.encode("t5 t17 t300") == [5, 17, 300]


def test_full_stack_is_built_with_the_swap_and_dummy_weights(tmp_path, group):
    S.install(sparse_block="ours")
    d = S.make_model_dir(str(tmp_path / "m"), **SMALL)
    from nanovllm.config import Config
    from nanovllm.models.glm4_moe import Glm4MoeForCausalLM
    from nanovllm.models.glm4_moe.mlp import Glm4MoeMLP as RefMLP
    hf = Config(d, max_num_batched_tokens=512, max_model_len=256).hf_config
    prev = torch.get_default_dtype()
    torch.set_default_dtype(hf.torch_dtype)
    try:
        model = Glm4MoeForCausalLM(hf)
    finally:
        torch.set_default_dtype(prev)
    layers = model.model.layers
    assert len(layers) == 3                                           # not the 2 of model.py:20-22
    assert isinstance(layers[0].mlp, RefMLP)                          # dense layer 0 (decode_layer.py:31-36)
    assert all(isinstance(l.mlp, glm4_moe_b200.Glm4MoeMoE) for l in layers[1:])
    model.load_weights(d)                                             # dummy_weights.json -> random init
    for name, p in model.named_parameters():
        assert bool(torch.isfinite(p.float()).all()), name
        if p.dim() == 2 and p.numel():
            assert 0.015 < float(p.float().std()) < 0.025, name
    assert float(layers[1].mlp.gate.e_score_correction_bias.max()) <= 0.1
    assert layers[1].mlp.gate.e_score_correction_bias.dtype == torch.float32
    # the two MoE layers draw different weights
    assert not torch.equal(layers[1].mlp._w_down, layers[2].mlp._w_down)
    # the residual reaches the final norm (model.py:75 as commented in the reference)
    class Stub(torch.nn.Module):
        def forward(self, positions, h, res):
            return torch.ones_like(h), (h if res is None else res + h)
    model.model.layers = torch.nn.ModuleList([Stub(), Stub()])
    ids = torch.tensor([3, 4, 5])
    out = model(ids, torch.arange(3))
    emb = model.model.embed_tokens(ids)
    want, _ = model.model.norm(torch.ones_like(emb), emb + torch.ones_like(emb))
    assert torch.equal(out, want)


def test_fp8_construction_keeps_no_model_dtype_stack(tmp_path, group):
    S.install(sparse_block="ours", weight_format="fp8")
    try:
        d = S.make_model_dir(str(tmp_path / "m"), **SMALL)
        from nanovllm.config import Config
        from nanovllm.models.glm4_moe import Glm4MoeForCausalLM
        hf = Config(d, max_num_batched_tokens=512, max_model_len=256).hf_config
        prev = torch.get_default_dtype()
        torch.set_default_dtype(hf.torch_dtype)
        try:
            model = Glm4MoeForCausalLM(hf)
        finally:
            torch.set_default_dtype(prev)
        model.load_weights(d)
        blk = model.model.layers[1].mlp
        assert blk.weight_format == "fp8" and blk._w_gate_up.numel() == 0
        assert blk._wq_gate_up.dtype == torch.float8_e4m3fn and blk._s_gate_up.shape == (9, 2, 2)
        deq = blk.dequantized_weights()["w_down"]
        assert 0.015 < float(deq.std()) < 0.025
    finally:
        S.install(sparse_block="ours", weight_format="native")
