"""Host-side shim that lets the UNMODIFIED reference engine run a full GLM4-MoE decoder stack with this
package's sparse block swapped in (BASELINE configs[3]).

The reference's GLM path is work in progress and cannot run a model as shipped.  Nothing
under `nanovllm/` is edited: `install()` imports the installed copy (`baseline/_ref/`, put there by
`oracle/make_ref.py`) and rebinds, at construction time only,

  decode_layer.py:6,29     `Glm4MoeMoE`        -> `glm4_moe_b200.Glm4MoeMoE` (the one-line swap of INTEGRATION.md)
  model.py:20-22           2 hard-coded layers  -> `config.num_hidden_layers` (the reference's own
                                                   commented-out line, model.py:19)
  model.py:75-76           final norm without the residual -> `self.norm(hidden_states, residual)`
                                                   (its commented-out line, model.py:75; cf. qwen3.py:176)
  moe.py:153-157, attention_new.py:288-292   per-forward debug `load_file` of a path that exists only on
                                                   the author's machine -> placeholder (oracle.reference_live
                                                   does the same for the checker)
  config.py:24 / attention_new.py:90-95       `hf_config.rope_theta`, `.rope_scaling`, `.head_dim`,
                                                   `.torch_dtype`: renamed / dropped by transformers 5.x
                                                   -> an attribute view that restores them
  model.py:107-132         `load_weights` from a checkpoint -> deterministic random init when the model
                                                   directory carries `dummy_weights.json` (no checkpoint and
                                                   no network in this environment)
  parallel_state.py:1-9    dummy TP group (world 1)  -> the real rank / world of the NCCL group the engine
                                                   creates (model_runner.py:27), so that with
                                                   `tensor_parallel_size = N` attention and the dense MLP are
                                                   tensor-parallel as the reference's own layers implement it
                                                   (linear.py, the KV cache already divides the heads,
                                                   model_runner.py:111) while the sparse blocks are
                                                   EXPERT-parallel (`expert_parallel="p2p" | "allreduce"`,
                                                   distributed.py): BASELINE configs[3] at 2/4/8 GPUs
  decode_layer.py:96-99    (opt-in, `fused_norm=True`) `self_attn` -> `post_attention_layernorm` -> `mlp` ->
                                                   the o_proj GEMM keeps its PARTIAL sums (its `tp_size` is set to 1
                                                   after construction, so `linear.py:153-154` skips the NCCL
                                                   all-reduce) and ONE library call does all-reduce + residual add
                                                   + RMSNorm + sparse block (`distributed.PeerComm`,
                                                   `csrc/ep_allreduce.cu`); `RMSNorm`
                                                   (decode_layer.py:8) becomes the one-launch kernel of
                                                   `layernorm.py` (the reference's is ~8 torch launches)
  llm_engine.py:26         `ModelRunner` as the spawn target -> `shim_model_runner` (same arguments), which
                                                   re-applies this shim in the spawned rank before constructing
                                                   the reference's `ModelRunner`

Attention (`nanovllm/layers/attention.py`, flash-attn), the KV cache, the scheduler, the sampler and the
CUDA-graph runner (`engine/model_runner.py:216-258`) run as shipped.  `make_model_dir()` writes the synthetic
model directory (`config.json`, a word-level tokenizer, `dummy_weights.json`) that `Config` / `LLMEngine`
insist on (`config.py:21-24`, `llm_engine.py:31`).
"""
from __future__ import annotations

import json
import os
import sys
import types

import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CANDIDATES = (os.path.join(ROOT, "baseline", "_ref"), "/root/reference")
DUMMY_MARKER = "dummy_weights.json"


def reference_root() -> str | None:
    for c in REF_CANDIDATES:
        if os.path.exists(os.path.join(c, "nanovllm", "engine", "model_runner.py")):
            return c
    return None


class HfConfigView:
    """Attribute view over a transformers config restoring the names the reference reads
    (`config.py:24`, `model_runner.py:29,108-112`, `attention_new.py:29,84-95`)."""

    def __init__(self, hf, dtype=None):
        object.__setattr__(self, "_hf", hf)
        rp = getattr(hf, "rope_parameters", None) or {}
        over = {}
        if "rope_theta" not in hf.__dict__:
            over["rope_theta"] = float(rp.get("rope_theta", 10000.0))
        # the reference asserts `rope_scaling is None` (rotary_embedding.py:66); transformers 5.x aliases the
        # attribute to the `rope_parameters` dict even for the default rope type
        over["rope_scaling"] = None if rp.get("rope_type", "default") == "default" else rp
        if "head_dim" not in hf.__dict__:
            over["head_dim"] = hf.hidden_size // hf.num_attention_heads
        td = dtype or hf.__dict__.get("dtype") or hf.__dict__.get("torch_dtype") or torch.bfloat16
        if isinstance(td, str):
            td = getattr(torch, td.replace("torch.", ""))
        over["torch_dtype"] = td
        object.__setattr__(self, "_over", over)

    def __getattr__(self, name):
        over = object.__getattribute__(self, "_over")
        if name in over:
            return over[name]
        return getattr(object.__getattribute__(self, "_hf"), name)

    def __setattr__(self, name, value):
        object.__getattribute__(self, "_over")[name] = value


_installed = {}
ENV_ARGS = "GLM4MOE_SHIM_ARGS"


class _TPGroup:
    """What `get_tp_group()` callers read (`linear.py:25-26`, `attention_new.py:16,117-119`): rank and
    world size of the default process group the engine initialises (`model_runner.py:27`)."""

    @property
    def world_size(self):
        import torch.distributed as dist
        return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1

    @property
    def rank(self):
        import torch.distributed as dist
        return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def shim_model_runner(config, rank, event):
    """Spawn target standing in for `ModelRunner` (llm_engine.py:26): a spawned rank starts from a fresh
    interpreter, so the rebinding is re-applied from the arguments the parent recorded, then the
    reference's own `ModelRunner` runs as shipped."""
    kw = json.loads(os.environ.get(ENV_ARGS, "{}"))
    if "runner" not in _installed:
        install(**kw)
    _installed["max_tokens"] = int(config.max_num_batched_tokens)
    if rank > 0:
        # a spawned rank blocks in ModelRunner.loop() for its whole life: leave when the parent is gone
        import threading
        import time
        parent = os.getppid()

        def watchdog():
            while os.getppid() == parent:
                time.sleep(1.0)
            os._exit(1)

        threading.Thread(target=watchdog, daemon=True).start()
    import torch.distributed as dist
    if not getattr(dist.init_process_group, "_glm4moe_patched", False):
        real_init = dist.init_process_group

        def init_process_group(backend=None, init_method=None, *a, **k):     # model_runner.py:27
            if isinstance(init_method, str):      # the container's "localhost" may not resolve
                init_method = init_method.replace("tcp://localhost:", "tcp://127.0.0.1:")
            return real_init(backend, init_method, *a, **k)

        init_process_group._glm4moe_patched = True
        dist.init_process_group = init_process_group
    return _installed["runner"](config, rank, event)


def install(sparse_block: str = "ours", weight_format: str = "native", share_workspace: bool = True,
            expert_parallel: str | None = None, dense_mlp: str = "reference", fused_norm: bool = False):
    """Import the reference package and apply the construction-time rebinding described above.
    `sparse_block`: "ours" (this package's `Glm4MoeMoE`) or "reference" (the reference's own block, for
    A/B runs of the same stack).  `expert_parallel`: layout of the sparse blocks when the engine runs
    `tensor_parallel_size > 1` ranks ("p2p": peer-memory combine, "allreduce": NCCL; None: every rank holds
    all experts).  `dense_mlp`: "ours" also swaps the dense MLP of the first layers (decode_layer.py:31-36) for
    `Glm4MoeDenseMLP` (the same tensor-core grouped GEMM).  `fused_norm`: the norms of the decoder
    layer run as this package's kernel and the post-attention norm (with the o_proj all-reduce when
    `tensor_parallel_size > 1`) is fused to the block that follows it.  Returns the `nanovllm` package."""
    os.environ[ENV_ARGS] = json.dumps(dict(sparse_block=sparse_block, weight_format=weight_format,
                                           share_workspace=share_workspace, expert_parallel=expert_parallel,
                                           dense_mlp=dense_mlp, fused_norm=bool(fused_norm)))
    _installed["fused_norm"] = bool(fused_norm) and sparse_block == "ours"
    _installed["ep_layout"] = expert_parallel if sparse_block == "ours" else None
    root = reference_root()
    if root is None:
        raise RuntimeError("the reference engine is not installed: run `python oracle/make_ref.py` where "
                           "/root/reference is mounted (baseline/_ref/ then travels with the repo)")
    if root not in sys.path:
        sys.path.insert(0, root)
    # debug file reads (moe.py:153-157, attention_new.py:288-292)
    import safetensors.torch as st
    if not getattr(st.load_file, "_glm4moe_patched", False):
        real = st.load_file

        def load_file(filename, *a, **k):
            if str(filename).startswith("/data/ai_infra/debug"):
                z = torch.zeros(1)
                return {"experts_output": z, "shared_output": z, "final_hidden_states": z}
            return real(filename, *a, **k)

        load_file._glm4moe_patched = True
        st.load_file = load_file

    import nanovllm
    from nanovllm import config as ref_config
    from nanovllm.models.glm4_moe import decode_layer as DL
    from nanovllm.models.glm4_moe import model as M
    from nanovllm.models.glm4_moe import moe as ref_moe
    from nanovllm.layers.embed_head import VocabParallelEmbedding
    from nanovllm.layers.layernorm import RMSNorm

    from nanovllm.engine import llm_engine as LE
    from nanovllm.engine import model_runner as MR
    if "ref_block" not in _installed:
        _installed["ref_dense"] = DL.Glm4MoeMLP
        _installed["ref_block"] = ref_moe.Glm4MoeMoE
        _installed["post_init"] = ref_config.Config.__post_init__
        _installed["runner"] = MR.ModelRunner
        _installed["ref_norm"] = DL.RMSNorm
        _installed["ref_layer_forward"] = DL.Glm4MoeDecoderLayer.forward
    LE.ModelRunner = shim_model_runner                       # llm_engine.py:26,30
    # parallel_state.py:1-9: every module that imported the dummy group by name
    tp = _TPGroup()
    for name, mod in list(sys.modules.items()):
        if name.startswith("nanovllm") and mod is not None and hasattr(mod, "get_tp_group"):
            mod.get_tp_group = lambda: tp

    from . import moe as ours
    if sparse_block == "ours":
        ours.Glm4MoeMoE.default_weight_format = weight_format
        ours.Glm4MoeMoE.share_workspace = bool(share_workspace)
        DL.Glm4MoeMoE = ours.Glm4MoeMoE                      # decode_layer.py:6 — the swap
    elif sparse_block == "reference":
        DL.Glm4MoeMoE = _installed["ref_block"]
    else:
        raise ValueError(sparse_block)
    if dense_mlp not in ("ours", "reference"):
        raise ValueError(dense_mlp)
    DL.Glm4MoeMLP = ours.Glm4MoeDenseMLP if dense_mlp == "ours" else _installed["ref_dense"]   # decode_layer.py:7,31

    from . import layernorm as our_norm
    Norm = our_norm.RMSNorm if _installed["fused_norm"] else _installed["ref_norm"]
    DL.RMSNorm = Norm                                        # decode_layer.py:8,37-38

    def layer_forward(self, positions, hidden_states, residual):     # decode_layer.py:80-100
        if residual is None:
            hidden_states, residual = self.input_layernorm(hidden_states), hidden_states
        else:
            hidden_states, residual = self.input_layernorm(hidden_states, residual)
        hidden_states = self.self_attn(hidden_states, positions)      # partial sums when the o_proj all-reduce is off
        fused = self.__dict__.get("_fused_post")
        if fused is not None:                                # all-reduce + add + norm + block: one library call
            return fused(hidden_states, residual)
        hidden_states, residual = self.post_attention_layernorm(hidden_states, residual)
        return self.mlp(hidden_states), residual

    DL.Glm4MoeDecoderLayer.forward = layer_forward if _installed["fused_norm"] else _installed["ref_layer_forward"]

    def model_init(self, config):                            # model.py:14-24 with the line of model.py:19
        nn.Module.__init__(self)
        import torch.distributed as dist
        world = dist.get_world_size() if dist.is_initialized() else 1
        ours.Glm4MoeMoE.default_ep = None
        ours.Glm4MoeMoE.default_max_tokens = _installed.get("max_tokens")
        if world > 1 and _installed.get("ep_layout"):
            layout = _installed["ep_layout"]
            ours.Glm4MoeMoE.default_ep = (dist.get_rank(), world, layout == "p2p")
            # layout "p2p": with share_workspace the blocks share ONE peer-mapped workspace and its mappings
            # (moe.py `_peer_workspace_pool`); a private one per layer is 3 GB x 45 at the reference's default
            # max_num_batched_tokens = 16384
        self.embed_tokens = VocabParallelEmbedding(config.vocab_size, config.hidden_size)
        self.layers = nn.ModuleList([DL.Glm4MoeDecoderLayer(config, prefix=f"model.layers.{i}")
                                     for i in range(config.num_hidden_layers)])
        self.norm = Norm(config.hidden_size, eps=config.rms_norm_eps)

    def model_forward(self, input_ids, positions):           # model.py:63-78 with the line of model.py:75
        hidden_states = self.embed_tokens(input_ids)
        residual = None
        for layer in self.layers:
            hidden_states, residual = layer(positions, hidden_states, residual)
        hidden_states, _ = self.norm(hidden_states, residual)
        return hidden_states

    M.Glm4MoeModel.__init__ = model_init
    M.Glm4MoeModel.forward = model_forward

    real_load = _installed.setdefault("load_weights", M.Glm4MoeForCausalLM.load_weights)

    def load_weights(self, model_path):                      # model.py:107-132
        marker = os.path.join(model_path, DUMMY_MARKER)
        if os.path.exists(marker):
            init_dummy_weights(self, **json.load(open(marker)))
        else:
            real_load(self, model_path)
        setup_expert_parallel(self)
        setup_fused_norm(self)

    M.Glm4MoeForCausalLM.load_weights = load_weights

    post_init = _installed["post_init"]

    def config_post_init(self):                              # config.py:20-26
        post_init(self)
        if not isinstance(self.hf_config, HfConfigView):
            self.hf_config = HfConfigView(self.hf_config)

    ref_config.Config.__post_init__ = config_post_init
    return nanovllm


def setup_expert_parallel(model) -> int:
    """Wrap every sparse block of an expert-parallel rank in `distributed.ExpertParallelMoE` (collective:
    every rank calls it at the same point, right after its weights exist).  Layout "p2p" allocates and
    exchanges the peer-mapped workspaces here, sized for the engine's largest batch, so nothing is
    allocated during warm-up or CUDA-graph capture.  Returns the number of blocks wrapped."""
    import torch.distributed as dist
    from . import moe as ours
    from .distributed import ExpertParallelMoE
    layout = _installed.get("ep_layout")
    if not layout or not dist.is_initialized() or dist.get_world_size() == 1:
        return 0
    n = 0
    for layer in model.model.layers:
        blk = layer.mlp
        if isinstance(blk, ours.Glm4MoeMoE) and blk.ep_size > 1:
            ep = ExpertParallelMoE(blk, blk.config.n_routed_experts, layout=layout)
            if layout == "p2p":
                ep.setup_p2p(_installed.get("max_tokens"))
            layer.mlp = ep
            n += 1
    return n


def setup_fused_norm(model) -> int:
    """`fused_norm=True`: give every decoder layer the one-call form of decode_layer.py:96-99 (collective when the
    engine runs several ranks: the peer-mapped communication buffer is allocated and exchanged here, sized for
    the engine's largest batch).  Returns the number of layers whose post-attention step is fused."""
    import torch.distributed as dist
    from . import moe as ours
    from .distributed import ExpertParallelMoE, PeerComm
    if not _installed.get("fused_norm"):
        return 0
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    layers = list(model.model.layers)
    comm = None
    if world > 1:
        w = layers[0].post_attention_layernorm.weight
        comm = PeerComm(_installed.get("max_tokens") or 4096, w.numel(), w.dtype, device=w.device)
        _installed["comm"] = comm
    return sum(fuse_layer(layer, world, comm) for layer in layers)


def fuse_layer(layer, world: int = 1, comm=None) -> bool:
    """Give ONE `Glm4MoeDecoderLayer` the one-call form of decode_layer.py:96-99 (the patched layer forward of
    `install(fused_norm=True)` routes through `_fused_post` when a layer has it).  Returns whether it was fused."""
    from . import moe as ours
    from .distributed import ExpertParallelMoE
    norm, mlp = layer.post_attention_layernorm, layer.mlp
    if world > 1:
        # RowParallelLinear.forward reads `tp_size` only to decide on the all-reduce (linear.py:153): the
        # weight shard was cut at construction, so the GEMM now returns this rank's partial sums
        layer.self_attn.o_proj.tp_size = 1
    if isinstance(mlp, ExpertParallelMoE):
        mlp.comm = comm
        fused = lambda h, r, m=mlp, nm=norm: m.forward_prenorm(h, r, nm.weight, nm.eps)
    elif isinstance(mlp, ours.Glm4MoeMoE) and world == 1:
        fused = lambda h, r, m=mlp, nm=norm: m.forward_prenorm(h, r, nm.weight, nm.eps)
    elif world > 1:
        def fused(h, r, m=mlp, nm=norm, c=comm):         # dense layer: fused all-reduce + add + norm, then the MLP
            x, r2 = c.allreduce_add_rmsnorm(h, r, nm.weight, nm.eps)
            return m(x), r2
    else:
        return False
    layer.__dict__["_fused_post"] = fused
    return True


def init_dummy_weights(model, seed: int = 1234, std: float = 0.02, **_):
    """Deterministic random init of a `Glm4MoeForCausalLM`: N(0, std^2) for every matrix, ones for the norm
    weights, U(0, 0.1) for the fp32 router bias; this package's blocks fill themselves
    (`Glm4MoeMoE.init_random_`, which also quantises into the e4m3 stacks of an FP8 block)."""
    from . import moe as ours
    own = set()
    with torch.no_grad():
        for li, layer in enumerate(model.model.layers):
            if isinstance(layer.mlp, ours.Glm4MoeMoE):
                layer.mlp.init_random_(seed * 1000 + li, std)
                own.update(id(t) for t in layer.mlp.parameters())
                own.update(id(t) for t in layer.mlp.buffers())
        sharded = {}
        for mname, mod in model.named_modules():
            wt = mod._parameters.get("weight")
            if wt is not None and hasattr(wt, "weight_loader") and hasattr(mod, "tp_size"):
                sharded[id(wt)] = mod
        for name, p in model.named_parameters():
            if id(p) in own or p.numel() == 0:
                continue
            if p.dim() == 1:
                p.fill_(1.0 if "norm" in name else 0.0)
            elif id(p) in sharded:
                _init_sharded_matrix(sharded[id(p)], p, name, seed, std)
            else:
                p.copy_(_draw(p.shape, name, seed, std, p.device).to(p.dtype))
        for name, b in model.named_buffers():
            if id(b) not in own and name.endswith("e_score_correction_bias"):
                g = torch.Generator(device=b.device).manual_seed(_name_seed(name, seed))
                b.copy_(torch.rand(b.shape, generator=g, device=b.device, dtype=torch.float32) * 0.1)


def _name_seed(name: str, seed: int) -> int:
    import zlib
    return (zlib.crc32(name.encode()) ^ (int(seed) * 0x9E3779B1)) & 0x7FFFFFFFFFFFFFFF


def _draw(shape, name: str, seed: int, std: float, device) -> torch.Tensor:
    """N(0, std^2) fp32 tensor that depends only on (parameter name, seed, shape): the same on every rank."""
    g = torch.Generator(device=device).manual_seed(_name_seed(name, seed))
    return torch.randn(tuple(shape), generator=g, device=device, dtype=torch.float32) * std


def _init_sharded_matrix(mod, p, name: str, seed: int, std: float) -> None:
    """Fill the tensor-parallel shard `p` of `mod.weight` the way a checkpoint would: draw the FULL matrix (every
    rank draws the same one, keyed by the parameter name) and hand it to the reference's own `weight_loader`,
    which narrows it to this rank (linear.py:67-72,89-95,116-130,144-149, embed_head.py:27-32).  The tpN model is
    then the tp1 model cut into shards, so token ids and digests compare across tensor-parallel sizes."""
    tp = int(mod.tp_size)
    loader = p.weight_loader
    cols = p.shape[1]
    if hasattr(mod, "output_sizes"):                               # MergedColumnParallelLinear: [gate; up]
        for i, rows in enumerate(mod.output_sizes):
            loader(p, _draw((rows, cols), f"{name}#{i}", seed, std, p.device).to(p.dtype), i)
    elif hasattr(mod, "num_kv_heads") and hasattr(mod, "head_size"):   # QKVParallelLinear: [q; k; v]
        for sid, heads in (("q", mod.num_heads), ("k", mod.num_kv_heads), ("v", mod.num_kv_heads)):
            loader(p, _draw((heads * tp * mod.head_size, cols), f"{name}#{sid}", seed, std, p.device).to(p.dtype), sid)
    else:                                                          # column / row / replicated linear, embeddings
        tp_dim = getattr(mod, "tp_dim", 0)                         # embeddings: vocabulary rows (embed_head.py:27-32)
        full = list(p.shape)
        if tp_dim is not None:
            full[tp_dim] *= tp
        loader(p, _draw(full, name, seed, std, p.device).to(p.dtype))


def make_model_dir(path: str, *, hidden_size: int, n_routed_experts: int, num_experts_per_tok: int,
                   moe_intermediate_size: int, intermediate_size: int, num_hidden_layers: int,
                   num_attention_heads: int, num_key_value_heads: int, head_dim: int, vocab_size: int,
                   routed_scaling_factor: float = 1.0, first_k_dense_replace: int = 1,
                   max_position_embeddings: int = 4096, dtype: str = "bfloat16", seed: int = 1234) -> str:
    """Synthetic GLM4-MoE model directory: what `Config` (config.py:21-24: a directory with a
    `config.json` AutoConfig can read) and `LLMEngine` (llm_engine.py:31: a fast tokenizer) need."""
    os.makedirs(path, exist_ok=True)
    cfg = {
        "architectures": ["Glm4MoeForCausalLM"], "model_type": "glm4_moe", "hidden_size": hidden_size,
        "n_routed_experts": n_routed_experts, "num_experts_per_tok": num_experts_per_tok,
        "moe_intermediate_size": moe_intermediate_size, "intermediate_size": intermediate_size,
        "n_shared_experts": 1, "n_group": 1, "topk_group": 1, "norm_topk_prob": True,
        "routed_scaling_factor": routed_scaling_factor, "first_k_dense_replace": first_k_dense_replace,
        "num_hidden_layers": num_hidden_layers, "num_attention_heads": num_attention_heads,
        "num_key_value_heads": num_key_value_heads, "head_dim": head_dim, "vocab_size": vocab_size,
        "max_position_embeddings": max_position_embeddings, "rms_norm_eps": 1e-5, "hidden_act": "silu",
        "attention_bias": False, "use_qk_norm": False, "partial_rotary_factor": 0.5, "rope_theta": 10000.0,
        "tie_word_embeddings": False, "dtype": dtype, "torch_dtype": dtype, "eos_token_id": 1, "pad_token_id": 0,
    }
    json.dump(cfg, open(os.path.join(path, "config.json"), "w"), indent=1)
    json.dump({"seed": seed, "std": 0.02}, open(os.path.join(path, DUMMY_MARKER), "w"))
    from tokenizers import Tokenizer, models, pre_tokenizers
    vocab = {"<pad>": 0, "</s>": 1, "<unk>": 2}
    for i in range(3, vocab_size):
        vocab[f"t{i}"] = i
    tok = Tokenizer(models.WordLevel(vocab=vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    tok.save(os.path.join(path, "tokenizer.json"))
    json.dump({"tokenizer_class": "PreTrainedTokenizerFast", "eos_token": "</s>", "pad_token": "<pad>",
               "unk_token": "<unk>", "model_max_length": max_position_embeddings},
              open(os.path.join(path, "tokenizer_config.json"), "w"))
    return path


AIR_MODEL = dict(hidden_size=4096, n_routed_experts=128, num_experts_per_tok=8, moe_intermediate_size=1408,
                 intermediate_size=10944, num_attention_heads=96, num_key_value_heads=8, head_dim=128,
                 vocab_size=151552, routed_scaling_factor=1.0, first_k_dense_replace=1)
