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


def test_multi_rank_plumbing(tmp_path):
    """What the shim adds for `tensor_parallel_size > 1` (BASELINE configs[3] at 2/4/8 GPUs): the dummy TP
    group becomes the real one in every module that imported it, the spawn target is picklable by name and
    carries the install arguments through the environment, the config view survives the spawn pickling."""
    import json
    import pickle
    S.install(sparse_block="ours", weight_format="fp8", expert_parallel="p2p")
    import nanovllm.layers.linear as lin
    from nanovllm.models.glm4_moe import attention_new
    from nanovllm.engine import llm_engine
    assert isinstance(lin.get_tp_group(), S._TPGroup) and isinstance(attention_new.get_tp_group(), S._TPGroup)
    g = lin.get_tp_group()
    assert (g.rank, g.world_size) == ((dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1))
    assert llm_engine.ModelRunner is S.shim_model_runner
    assert pickle.loads(pickle.dumps(S.shim_model_runner)) is S.shim_model_runner
    assert json.loads(os.environ[S.ENV_ARGS]) == dict(sparse_block="ours", weight_format="fp8", share_workspace=True,
                                                      expert_parallel="p2p", dense_mlp="reference",
                                                      fused_norm=False)
    d = S.make_model_dir(str(tmp_path / "m"), **SMALL)
    from nanovllm.config import Config
    cfg = pickle.loads(pickle.dumps(Config(d, max_num_batched_tokens=512, max_model_len=256)))
    assert cfg.hf_config.head_dim == 64 and cfg.hf_config.rope_scaling is None
    assert cfg.hf_config.torch_dtype == torch.bfloat16
    S.install()            # back to the defaults for the other tests
    assert json.loads(os.environ[S.ENV_ARGS])["expert_parallel"] is None


def test_fused_norm_plumbing(tmp_path, group):
    """`install(fused_norm=True)` (SURVEY 8f-2 part i / 8f-4): decode_layer.py:8 resolves to this package's RMSNorm,
    every sparse layer carries the one-call form of decode_layer.py:98-99 and the patched layer forward routes
    through it (stub callable: the real one needs the GPU library), the dense layer keeps the two-step form on
    one rank; `install()` puts the reference's own forward and norm back."""
    S.install()
    from nanovllm.models.glm4_moe import decode_layer as DL
    ref_forward = S._installed.get("ref_layer_forward", DL.Glm4MoeDecoderLayer.forward)
    S.install(sparse_block="ours", fused_norm=True)
    try:
        from glm4_moe_b200 import layernorm as LN
        assert DL.RMSNorm is LN.RMSNorm and DL.Glm4MoeDecoderLayer.forward is not ref_forward
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
        layers = model.model.layers
        assert isinstance(layers[1].input_layernorm, LN.RMSNorm) and isinstance(model.model.norm, LN.RMSNorm)
        assert "_fused_post" not in layers[0].__dict__                   # dense layer, one rank: nothing to fuse
        assert all("_fused_post" in l.__dict__ for l in layers[1:])
        # the patched forward hands (attention output, residual) to the fused callable and returns its pair
        lay = layers[1]
        calls = []
        class Norm(torch.nn.Module):
            def forward(self, h, r=None):
                return (h + 1, h) if r is not None else h + 1

        class Attn(torch.nn.Module):
            def forward(self, h, pos):
                return h * 2

        lay.input_layernorm, lay.self_attn = Norm(), Attn()
        lay.__dict__["_fused_post"] = lambda h, r: (calls.append((h.clone(), r.clone())) or (h - r, r + 7))
        x = torch.ones(3, 4)
        out, res = lay(torch.arange(3), x, None)
        assert torch.equal(calls[0][0], (x + 1) * 2) and torch.equal(calls[0][1], x)
        assert torch.equal(out, (x + 1) * 2 - x) and torch.equal(res, x + 7)
    finally:
        S.install()
    assert DL.Glm4MoeDecoderLayer.forward is ref_forward and DL.RMSNorm is S._installed["ref_norm"]


def test_random_init_of_an_expert_parallel_shard_equals_the_single_gpu_block(tmp_path):
    """`init_random_` seeds every expert by its global id: rank r's shard holds experts [r*E/R, (r+1)*E/R) of
    the single-GPU block, with the same shared expert, gate and bias (identical routing on every rank)."""
    S.install()
    d = S.make_model_dir(str(tmp_path / "m"), **SMALL)
    from nanovllm.config import Config
    hf = Config(d, max_num_batched_tokens=512, max_model_len=256).hf_config
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        one = glm4_moe_b200.Glm4MoeMoE(hf).init_random_(7)
        glm4_moe_b200.Glm4MoeMoE.default_ep = (1, 2, False)
        try:
            shard = glm4_moe_b200.Glm4MoeMoE(hf)          # as decode_layer.py:29 constructs it
        finally:
            glm4_moe_b200.Glm4MoeMoE.default_ep = None
        shard.init_random_(7)
    finally:
        torch.set_default_dtype(prev)
    E = hf.n_routed_experts
    assert (shard.ep_rank, shard.ep_size, shard.expert_begin, shard.expert_end) == (1, 2, E // 2, E)
    assert torch.equal(shard._w_gate_up[:-1], one._w_gate_up[E // 2:E])
    assert torch.equal(shard._w_down[:-1], one._w_down[E // 2:E])
    assert torch.equal(shard._w_gate_up[-1], one._w_gate_up[-1]) and torch.equal(shard._w_down[-1], one._w_down[-1])
    assert torch.equal(shard.gate.weight, one.gate.weight)
    assert torch.equal(shard.gate.e_score_correction_bias, one.gate.e_score_correction_bias)
    assert not torch.equal(one._w_gate_up[0], one._w_gate_up[1])


def test_dummy_weights_of_a_tensor_parallel_rank_are_shards_of_the_tp1_model(monkeypatch):
    """`init_dummy_weights` draws every matrix in full, keyed by its parameter name, and lets the reference's own
    `weight_loader` cut this rank's shard: concatenating what ranks 0 and 1 of a tp2 run hold gives the tp1 tensor,
    for every linear flavour the decoder uses (qkv, merged gate/up, row-parallel, column-parallel, embedding)."""
    from torch import nn
    import nanovllm.layers.linear as lin
    import nanovllm.layers.embed_head as eh

    class FakeGroup:
        def __init__(self, rank, world):
            self.rank, self.world_size = rank, world

    def build(rank, world):
        monkeypatch.setattr(lin, "get_tp_group", lambda: FakeGroup(rank, world))
        monkeypatch.setattr(eh.dist, "get_rank", lambda *a, **k: rank)
        monkeypatch.setattr(eh.dist, "get_world_size", lambda *a, **k: world)

        class Net(nn.Module):
            def __init__(self):
                super().__init__()
                self.qkv = lin.QKVParallelLinear(64, 16, 4, 2)
                self.gate_up = lin.MergedColumnParallelLinear(64, [96, 96])
                self.down = lin.RowParallelLinear(96, 64)
                self.col = lin.ColumnParallelLinear(64, 32)
                self.rep = lin.ReplicatedLinear(64, 8)
                self.emb = eh.VocabParallelEmbedding(128, 64)
                self.norm = nn.LayerNorm(64)

        class Model(nn.Module):
            def __init__(self):
                super().__init__()
                self.model = nn.Module()
                self.model.layers = nn.ModuleList()
                self.net = Net()

        m = Model()
        S.init_dummy_weights(m, seed=7, std=0.02)
        return m.net

    one, r0, r1 = build(0, 1), build(0, 2), build(1, 2)
    q, kv = 4 * 16, 2 * 16
    full = one.qkv.weight
    for lo, n in ((0, q), (q, kv), (q + kv, kv)):               # q, k, v blocks: each cut in half by heads
        h = n // 2
        off = {0: 0, q: q // 2, q + kv: q // 2 + kv // 2}[lo]
        assert torch.equal(torch.cat([r0.qkv.weight[off:off + h], r1.qkv.weight[off:off + h]]), full[lo:lo + n])
    for i in range(2):                                          # gate rows, up rows
        got = torch.cat([r0.gate_up.weight[i * 48:(i + 1) * 48], r1.gate_up.weight[i * 48:(i + 1) * 48]])
        assert torch.equal(got, one.gate_up.weight[i * 96:(i + 1) * 96])
    assert torch.equal(torch.cat([r0.down.weight, r1.down.weight], dim=1), one.down.weight)
    assert torch.equal(torch.cat([r0.col.weight, r1.col.weight], dim=0), one.col.weight)
    assert torch.equal(torch.cat([r0.emb.weight, r1.emb.weight], dim=0), one.emb.weight)
    assert torch.equal(r0.rep.weight, one.rep.weight) and torch.equal(r1.rep.weight, one.rep.weight)
    assert not torch.equal(r0.col.weight, r1.col.weight)        # the shards are different slices, not one tensor twice
    assert 0.015 < float(one.qkv.weight.std()) < 0.025 and bool((one.norm.weight == 1).all())
