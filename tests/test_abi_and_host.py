"""CPU: the C-ABI library builds, loads and exports every symbol include/glm4_moe_b200.h declares;
host-side module logic (parameter names, stack aliasing, loader, loud failure without CUDA)."""
import os
import re

import pytest
import torch

import glm4_moe_b200
from glm4_moe_b200 import synth
from glm4_moe_b200 import _lib
from glm4_moe_b200.build import build_library, LIB_PATH
from conftest import ROOT


@pytest.fixture(scope="session", autouse=True)
def built_library():
    build_library()
    assert os.path.exists(LIB_PATH)


def _cfg(shape):
    from transformers.models.glm4_moe import Glm4MoeConfig
    return Glm4MoeConfig(hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                         moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                         routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)


def test_header_symbols_are_exported():
    hdr = open(os.path.join(ROOT, "include", "glm4_moe_b200.h")).read()
    declared = set(re.findall(r"\b(glm4moe_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.glm4moe_abi_version() == 2 == _lib.ABI_VERSION
    assert lib.glm4moe_error_string(-2).decode().startswith("unsupported")


def test_no_torch_types_in_abi():
    hdr = open(os.path.join(ROOT, "include", "glm4_moe_b200.h")).read()
    assert "torch" not in hdr and "at::" not in hdr and "Tensor" not in hdr


def test_header_is_plain_c99(tmp_path):
    """The boundary is a C ABI: the header must compile as C99 on its own (no C++, no torch, no CUDA headers) and a C
    host must link against the shared library with nothing but the header."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    src = tmp_path / "host.c"
    src.write_text('#include "glm4_moe_b200.h"\n#include <stdio.h>\n'
                   'int main(void) {\n'
                   '  glm4moe_config_t c = {4096, 128, 8, 1408, 1.0f, GLM4MOE_BF16, GLM4MOE_W_NATIVE, 256, 0, 128, 1, 0};\n'
                   '  printf("%d %zu %zu\\n", glm4moe_abi_version(), glm4moe_workspace_bytes(&c),\n'
                   '         glm4moe_ep_comm_bytes(256, 4096, GLM4MOE_BF16, 8));\n'
                   '  return 0;\n}\n')
    inc = os.path.join(ROOT, "include")
    subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-I", inc, str(src)], check=True)
    from glm4_moe_b200.build import LIB_PATH
    exe = tmp_path / "host"
    subprocess.run([gcc, "-std=c99", "-I", inc, str(src), "-o", str(exe), LIB_PATH,
                    "-Wl,-rpath," + os.path.dirname(LIB_PATH)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    lib = _lib.load()
    assert int(out[0]) == 2 and int(out[2]) == lib.glm4moe_ep_comm_bytes(256, 4096, _lib.BF16, 8) and int(out[1]) > 0


def test_workspace_and_arg_errors_without_gpu():
    lib = _lib.load()
    import ctypes as C
    s = synth.AIR
    cfg = _lib.Config(s.H, s.E, s.k, s.I, 1.0, _lib.BF16, _lib.W_NATIVE, 256, 0, s.E, 1, 0)
    n = lib.glm4moe_workspace_bytes(C.byref(cfg))
    rows = 256 * (s.k + 1)
    assert n > rows * (2 * s.H + s.I) * 2          # x_rows + y + act at least
    bad = _lib.Config(s.H, 300, s.k, s.I, 1.0, _lib.BF16, 0, 256, 0, 300, 1, 0)   # E > 256
    assert lib.glm4moe_workspace_bytes(C.byref(bad)) == 0
    assert lib.glm4moe_router_logits(None, None, None, 1, 8, 8, _lib.BF16, None) == -1
    assert lib.glm4moe_tile_capacity(32, 8, 128, 16) == 128 + 16 + 2 + 1


def test_front_end_entries_without_gpu():
    """Host arithmetic and argument checks of the fused front end (no compute call without a GPU): the
    communication buffer of the peer-memory all-reduce (csrc/ep_allreduce.cuh: 1 KB of flags, 2 x R one-shot
    inboxes, 2 x R reduce-scatter inboxes, 2 all-gather inboxes, R x 1024 row flags), refusals, NULL checks."""
    import ctypes as C
    lib = _lib.load()
    H, es = 4096, 2
    for T, R in ((32, 2), (8192, 8), (100, 3)):
        n = lib.glm4moe_ep_comm_bytes(T, H, _lib.BF16, R)
        full = (T * H * es + 1023) // 1024 * 1024
        slot = min(full, 1 << 20)
        per = (-(-T // R) * H * es + 1023) // 1024 * 1024
        assert n == 1024 + 2 * R * slot + 2 * R * per + 2 * full + (R * 1024 * 4 + 1023) // 1024 * 1024, (T, R, n)
    assert lib.glm4moe_ep_comm_bytes(32, H, _lib.F32, 2) == 0          # 2-byte dtypes only
    assert lib.glm4moe_ep_comm_bytes(32, H, _lib.BF16, 9) == 0         # one NVSwitch box
    assert lib.glm4moe_ep_comm_bytes(32, 4090, _lib.BF16, 2) == 0      # 16-byte rows
    assert lib.glm4moe_ep_comm_bytes(0, H, _lib.BF16, 2) == 0
    assert lib.glm4moe_ep_comm_bytes(32, 8192, _lib.BF16, 2) > 0
    assert lib.glm4moe_ep_comm_bytes(32, 8200, _lib.BF16, 2) == 0      # a row the all-reduce kernels cannot hold: refused at set-up
    bases = (C.c_void_p * 2)(None, None)
    assert lib.glm4moe_ep_allreduce_add_rmsnorm(None, 0, 2, 32, None, None, None, 1e-5, 4, H, _lib.BF16, None, None,
                                                None) == -1
    assert lib.glm4moe_ep_allreduce_add_rmsnorm(bases, 2, 2, 32, 8, None, None, 1e-5, 4, H, _lib.BF16, 8, None,
                                                None) == -1            # rank outside the world
    assert lib.glm4moe_ep_allreduce_add_rmsnorm(bases, 0, 2, 32, 8, None, None, 1e-5, 4, H, _lib.F32, 8, None,
                                                None) == -2            # fp32 activations: unsupported
    assert lib.glm4moe_ep_allreduce_add_rmsnorm(bases, 0, 2, 32, 8, None, None, 1e-5, 4, H, _lib.BF16, 8, None,
                                                None) == -1            # unmapped peer buffer
    assert lib.glm4moe_layer_forward_prenorm(None, None, None, None, 1e-5, None, None, 1, None) == -1
    assert lib.glm4moe_layer_forward_ep_p2p_prenorm(None, None, None, None, 1e-5, None, None, 1, None, None, 32, 0, 1,
                                                    None) == -1
    assert lib.glm4moe_debug_set(11, 1) == 0 and lib.glm4moe_debug_set(11, -1) == 0     # KNOB_EP_AR_MODE
    assert lib.glm4moe_debug_set(12, 2) == 0 and lib.glm4moe_debug_set(12, 0) == 0      # KNOB_SPEC_PREFETCH


def test_module_names_match_reference_and_alias_stacks():
    shape = synth.TINY
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
    names = [n for n, _ in moe.named_parameters()]
    assert len(names) == 2 * shape.E + 3          # 35 parameters at E=16
    assert "experts.0.gate_up_proj.weight" in names and "experts.15.down_proj.weight" in names
    assert "gate.weight" in names and "shared_experts.gate_up_proj.weight" in names
    assert [n for n, _ in moe.named_buffers()] == ["gate.e_score_correction_bias"]
    assert moe.gate.e_score_correction_bias.dtype == torch.float32
    assert moe.experts[3].gate_up_proj.weight.shape == (2 * shape.I, shape.H)
    assert moe.experts[3].down_proj.weight.shape == (shape.H, shape.I)
    # reference-style initialisation through the per-expert parameters lands in the stacks
    moe.experts[3].gate_up_proj.weight.data.fill_(0.5)
    moe.shared_experts.down_proj.weight.data.fill_(-2.0)
    assert float(moe._w_gate_up[3].min()) == 0.5 and float(moe._w_down[shape.E].max()) == -2.0
    # dtype conversion keeps the aliasing
    moe = moe.to(torch.bfloat16)
    moe.experts[1].down_proj.weight.data.fill_(1.25)
    assert moe._w_down.dtype == torch.bfloat16 and float(moe._w_down[1].min()) == 1.25


def test_forward_fails_loudly_on_cpu():
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(synth.MINI))
    with pytest.raises(_lib.Glm4MoeLibraryError):
        moe(torch.zeros(4, synth.MINI.H))
    with pytest.raises(_lib.Glm4MoeLibraryError):
        moe.gate(torch.zeros(4, synth.MINI.H))


def test_load_weights_hf_layout(tmp_path):
    """moe.py:71-126 / mlp.py:83-131 key layout, incl. the experts.1 vs experts.1x prefix trap."""
    from safetensors.torch import save_file
    shape = synth.MoeShape(64, 12, 2, 32, 1.0, "loader")
    w = synth.make_weights(shape, torch.float32)
    p = "model.layers.1.mlp"
    sd = {f"{p}.gate.weight": w["gate_weight"], f"{p}.gate.e_score_correction_bias": w["bias"]}
    for e in range(shape.E + 1):
        q = f"{p}.experts.{e}" if e < shape.E else f"{p}.shared_experts"
        sd[f"{q}.gate_proj.weight"] = w["w_gate_up"][e, :shape.I].contiguous()
        sd[f"{q}.up_proj.weight"] = w["w_gate_up"][e, shape.I:].contiguous()
        sd[f"{q}.down_proj.weight"] = w["w_down"][e].contiguous()
    sd["model.layers.1.self_attn.o_proj.weight"] = torch.zeros(4, 4)
    keys = sorted(sd)
    save_file({k: sd[k] for k in keys[::2]}, str(tmp_path / "a.safetensors"))
    save_file({k: sd[k] for k in keys[1::2]}, str(tmp_path / "b.safetensors"))
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
    moe.load_weights(str(tmp_path), p)
    assert torch.equal(moe._w_gate_up, w["w_gate_up"]) and torch.equal(moe._w_down, w["w_down"])
    assert torch.equal(moe.gate.weight.data, w["gate_weight"])
    assert torch.equal(moe.gate.e_score_correction_bias, w["bias"])
    assert torch.equal(moe.experts[1].down_proj.weight.data, w["w_down"][1])     # not experts.10/11
    with pytest.raises(ValueError):
        moe.load_weights(str(tmp_path), "model.layers.7.mlp")
    # expert-parallel shard reads only its experts
    half = glm4_moe_b200.Glm4MoeMoE(_cfg(shape), ep_rank=1, ep_size=2)
    half.load_weights(str(tmp_path), p)
    assert torch.equal(half._w_gate_up[:6], w["w_gate_up"][6:12])
    assert torch.equal(half._w_down[6], w["w_down"][12])


def test_load_weights_fp8_block_checkpoint(tmp_path):
    """An FP8 checkpoint (e4m3 `weight` + fp32 `weight_scale_inv [rows/128, cols/128]`) fills the quantised
    stacks byte for byte, concatenates gate/up scales along dim 0 like the weights (mlp.py:118), switches
    the block to the e4m3 kernels and dequantises back to the source within e4m3 rounding."""
    from safetensors.torch import save_file
    from glm4_moe_b200 import fp8
    shape = synth.MoeShape(256, 4, 2, 128, 1.0, "loader-fp8")
    w = synth.make_weights(shape, torch.float32)
    p = "model.layers.3.mlp"
    sd = {f"{p}.gate.weight": w["gate_weight"], f"{p}.gate.e_score_correction_bias": w["bias"]}
    qs = {}
    for e in range(shape.E + 1):
        q = f"{p}.experts.{e}" if e < shape.E else f"{p}.shared_experts"
        for proj, t in (("gate_proj", w["w_gate_up"][e, :shape.I]), ("up_proj", w["w_gate_up"][e, shape.I:]),
                        ("down_proj", w["w_down"][e])):
            tq, ts = fp8.quantize_block128(t.contiguous())
            sd[f"{q}.{proj}.weight"], sd[f"{q}.{proj}.weight_scale_inv"] = tq, ts
            qs[(e, proj)] = (tq, ts)
    save_file(sd, str(tmp_path / "m.safetensors"))
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape), ep_rank=1, ep_size=2)      # experts 2, 3 + shared
    moe.load_weights(str(tmp_path / "m.safetensors"), p)
    assert moe.weight_format == "fp8"
    for slot, e in enumerate((2, 3, shape.E)):
        g, u, d = qs[(e, "gate_proj")], qs[(e, "up_proj")], qs[(e, "down_proj")]
        assert torch.equal(moe._wq_gate_up[slot].view(torch.uint8), torch.cat([g[0], u[0]]).view(torch.uint8))
        assert torch.equal(moe._s_gate_up[slot], torch.cat([g[1], u[1]]))
        assert torch.equal(moe._wq_down[slot].view(torch.uint8), d[0].view(torch.uint8))
        assert torch.equal(moe._s_down[slot], d[1])
    deq = moe.dequantized_weights()
    src = w["w_gate_up"][[2, 3, shape.E]]
    assert float((deq["w_gate_up"] - src).abs().max()) <= float(src.abs().max()) * 2 ** -4


def test_offline_fp8_conversion_round_trip(tmp_path):
    """`fp8.convert_checkpoint`: a bf16 HF-layout checkpoint becomes an FP8 block checkpoint
    that `load_weights` accepts; the loaded stacks equal quantising the source stacks directly, and
    non-expert tensors pass through untouched."""
    from safetensors.torch import save_file, load_file
    from glm4_moe_b200 import fp8
    shape = synth.MoeShape(256, 4, 2, 128, 1.0, "convert-fp8")
    w = synth.make_weights(shape, torch.bfloat16)
    p = "model.layers.1.mlp"
    sd = {f"{p}.gate.weight": w["gate_weight"], f"{p}.gate.e_score_correction_bias": w["bias"],
          "model.layers.1.input_layernorm.weight": torch.ones(shape.H, dtype=torch.bfloat16)}
    for e in range(shape.E + 1):
        q = f"{p}.experts.{e}" if e < shape.E else f"{p}.shared_experts"
        sd[f"{q}.gate_proj.weight"] = w["w_gate_up"][e, :shape.I].contiguous()
        sd[f"{q}.up_proj.weight"] = w["w_gate_up"][e, shape.I:].contiguous()
        sd[f"{q}.down_proj.weight"] = w["w_down"][e].contiguous()
    os.makedirs(tmp_path / "src")
    save_file(sd, str(tmp_path / "src" / "model-00001.safetensors"))
    stats = fp8.convert_checkpoint(str(tmp_path / "src"), str(tmp_path / "dst"))
    assert stats == {"converted": 3 * (shape.E + 1), "copied": 3}
    out = load_file(str(tmp_path / "dst" / "model-00001.safetensors"))
    assert torch.equal(out["model.layers.1.input_layernorm.weight"], sd["model.layers.1.input_layernorm.weight"])
    assert out[f"{p}.gate.weight"].dtype == torch.bfloat16
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
    moe.load_weights(str(tmp_path / "dst"), p)
    assert moe.weight_format == "fp8"
    # gate/up are quantised per projection (their 128-row blocks never straddle the two)
    q_gu, s_gu = fp8.quantize_block128(w["w_gate_up"])
    q_d, s_d = fp8.quantize_block128(w["w_down"])
    assert torch.equal(moe._wq_gate_up.view(torch.uint8), q_gu.view(torch.uint8))
    assert torch.equal(moe._s_gate_up, s_gu)
    assert torch.equal(moe._wq_down.view(torch.uint8), q_d.view(torch.uint8))
    assert torch.equal(moe._s_down, s_d)
    assert not moe.fp8_pow2_scales
    # power-of-two (UE8M0) block scales: detected at load, selects GLM4MOE_W_FP8_BLOCK128_POW2
    fp8.convert_checkpoint(str(tmp_path / "src"), str(tmp_path / "dst2"), pow2_scales=True)
    moe2 = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
    moe2.load_weights(str(tmp_path / "dst2"), p)
    assert moe2.weight_format == "fp8" and moe2.fp8_pow2_scales
    assert moe2._lib_config(64).weight_format == _lib.W_FP8_BLOCK128_POW2
    q2, s2 = fp8.quantize_block128(w["w_down"], pow2_scales=True)
    assert torch.equal(moe2._s_down, s2) and fp8.is_pow2(s2)
    assert bool(((s2 >= s_d) & (s2 < 2 * s_d)).all())            # the smallest power of two >= amax/448
    assert torch.equal(moe2._wq_down.view(torch.uint8), q2.view(torch.uint8))


def test_pow2_scale_rounding_and_workspace():
    """UE8M0 rounding used by the quant tier's power-of-two variant (fp8.py, oracle, common.cuh pow2_scale):
    exact powers stay, everything else goes up; the oracle's activation round trip is exact on values that
    are e4m3-representable under such a scale; the POW2 layout carries the packed scale words on top."""
    from glm4_moe_b200 import fp8
    from oracle import moe_oracle as O
    s = torch.tensor([1.0, 0.75, 0.5, 0.500001, 3e-5, 2.0 ** -20, 1.9999], dtype=torch.float32)
    want = torch.tensor([1.0, 1.0, 0.5, 1.0, 2.0 ** -15, 2.0 ** -20, 2.0], dtype=torch.float32)
    assert torch.equal(fp8.pow2_ceil(s), want)
    assert fp8.is_pow2(want) and not fp8.is_pow2(s)
    g = torch.Generator().manual_seed(3)
    a = torch.randn(5, 256, generator=g)
    r = O.fp8_act_roundtrip(a, pow2_scales=True)
    assert torch.equal(O.fp8_act_roundtrip(r, pow2_scales=True), r)          # idempotent
    q, sc = fp8.quantize_rows_1x128(a, pow2_scales=True)
    assert fp8.is_pow2(sc) and torch.equal((q.float().reshape(5, 2, 128) * sc.unsqueeze(-1)).reshape(5, 256), r)
    assert float((r - a).abs().max()) <= float(a.abs().max()) * 2.0 ** -3      # e4m3: 3 mantissa bits, <= 2x headroom lost
    lib = _lib.load()
    import ctypes as C
    sh = synth.AIR
    sizes = [lib.glm4moe_workspace_bytes(C.byref(_lib.Config(sh.H, sh.E, sh.k, sh.I, 1.0, _lib.BF16, f, 256, 0,
                                                             sh.E, 1, 0))) for f in (1, 2, 3)]
    rows = 256 * (sh.k + 1)
    assert sizes[1] >= sizes[0] + rows * 4 * (8 + 3) and sizes[2] == 0


def test_casts_keep_bias_fp32_and_move_only_the_stacks():
    """`.bfloat16()` / `.to(dtype)` cast floating-point buffers (nn.Module._apply); the library
    reads `e_score_correction_bias` as E fp32 values, so it must stay fp32 with its exact values, and
    per-expert parameters must stay views of the two stacks (no second copy)."""
    shape = synth.MINI
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
    w = synth.make_weights(shape, torch.float32)
    moe.load_stacked(w)
    for cast in (lambda m: m.bfloat16(), lambda m: m.to(torch.float16), lambda m: m.to(torch.bfloat16)):
        m = cast(moe)
        assert m.gate.e_score_correction_bias.dtype == torch.float32
        assert torch.equal(m.gate.e_score_correction_bias, w["bias"])
        assert m.gate.weight.dtype in (torch.bfloat16, torch.float16)
        base = m._w_gate_up.untyped_storage().data_ptr()
        for i, mlp in enumerate(m._mlps()):
            assert mlp.gate_up_proj.weight.untyped_storage().data_ptr() == base
            assert mlp.gate_up_proj.weight.shape == (2 * shape.I, shape.H)
            assert mlp.down_proj.weight.data_ptr() == m._w_down[i].data_ptr()
    # a bias of the wrong dtype is rejected before its pointer reaches the library
    moe.gate.e_score_correction_bias = moe.gate.e_score_correction_bias.to(torch.bfloat16)
    with pytest.raises(_lib.Glm4MoeLibraryError):
        moe.gate.bias_ptr()


def test_apply_after_dropping_the_native_stacks():
    """`.to()` after `quantize_fp8(drop_native=True)` (empty native stacks) must not index them."""
    shape = synth.MoeShape(256, 4, 2, 128, 1.0, "drop-native")
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
    finally:
        torch.set_default_dtype(prev)
    moe.load_stacked(synth.make_weights(shape, torch.bfloat16))
    moe.quantize_fp8(drop_native=True)
    moe = moe.to(torch.bfloat16).to("cpu")
    assert moe._w_gate_up.numel() == 0 and moe.experts[0].gate_up_proj.weight.numel() == 0
    assert moe._wq_gate_up.shape == (shape.E + 1, 2 * shape.I, shape.H)
    assert moe.gate.e_score_correction_bias.dtype == torch.float32


def test_dense_mlp_mirrors_the_reference_class_and_shards_like_its_linear_layers(tmp_path):
    """`Glm4MoeDenseMLP`: the reference dense class's constructor and parameter names
    (mlp.py:12-47), `load_weights(path, prefix)` from the HF keys, tensor-parallel shards = the rows of gate / up
    and the columns of down that `MergedColumnParallelLinear` / `RowParallelLinear` give rank r
    (linear.py:79-155); CPU tensors are refused loudly."""
    from safetensors.torch import save_file
    H, I = 256, 200
    g = torch.Generator().manual_seed(5)
    w = {k_: (torch.randn(s, generator=g) * 0.05).to(torch.bfloat16)
         for k_, s in (("gate_proj", (I, H)), ("up_proj", (I, H)), ("down_proj", (H, I)))}
    p = "model.layers.0.mlp"
    save_file({f"{p}.{k_}.weight": v for k_, v in w.items()} | {f"{p}x.gate_proj.weight": w["gate_proj"] * 0},
              str(tmp_path / "m.safetensors"))
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        one = glm4_moe_b200.Glm4MoeDenseMLP(hidden_size=H, intermediate_size=I, hidden_act="silu", prefix=p)
        shards = [glm4_moe_b200.Glm4MoeDenseMLP(H, I, "silu", prefix=p, tp_rank=r, tp_size=2) for r in range(2)]
        with pytest.raises(ValueError):
            glm4_moe_b200.Glm4MoeDenseMLP(H, I, "gelu")
    finally:
        torch.set_default_dtype(prev)
    assert [n for n, _ in one.named_parameters()] == ["gate_up_proj.weight", "down_proj.weight"]
    assert one.gate_up_proj.weight.shape == (2 * I, H) and one.down_proj.weight.dtype == torch.bfloat16
    one.load_weights(str(tmp_path), p)
    assert torch.equal(one.gate_up_proj.weight, torch.cat([w["gate_proj"], w["up_proj"]]))
    assert torch.equal(one.down_proj.weight, w["down_proj"])
    for r, m in enumerate(shards):
        m.load_weights(str(tmp_path / "m.safetensors"))
        sl = slice(r * I // 2, (r + 1) * I // 2)
        assert torch.equal(m.gate_up_proj.weight, torch.cat([w["gate_proj"][sl], w["up_proj"][sl]]))
        assert torch.equal(m.down_proj.weight, w["down_proj"][:, sl])
    with pytest.raises(ValueError):
        one.load_weights(str(tmp_path), "model.layers.9.mlp")
    with pytest.raises(_lib.Glm4MoeLibraryError):
        one(torch.zeros(3, H, dtype=torch.bfloat16))
    lib = _lib.load()
    assert lib.glm4moe_mlp_forward(None, None, None, None, None, 4, 256, 200, _lib.BF16, None, None) == -1


def test_ep_error_record_entries_without_gpu():
    """No failure recorded -> 0 and an untouched info array; the timeout setter accepts any value (<= 0 = default)
    and neither entry needs a device."""
    import ctypes as C
    lib = _lib.load()
    info = (C.c_int32 * 6)(*([7] * 6))
    assert lib.glm4moe_ep_last_error(info, 1) == 0 and lib.glm4moe_ep_last_error(None, 0) == 0
    assert _lib.ep_last_error() is None
    assert lib.glm4moe_ep_set_timeout_ms(250) == 0 and lib.glm4moe_ep_set_timeout_ms(0) == 0
    assert b"timed out" in lib.glm4moe_error_string(-5) and b"disagree" in lib.glm4moe_error_string(-6)
