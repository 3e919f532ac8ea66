"""Drop-in replacement for the reference's GLM4-MoE sparse block.

Mirrors `nanovllm/models/glm4_moe/moe.py` (`GlmMoeSelectTopk` :13-50, `Glm4MoeMoE` :52-176) and
`nanovllm/models/glm4_moe/mlp.py` (`Glm4MoeMLP` :12-54): same constructor argument, same
`forward(hidden_states)`, same `load_weights(path, prefix)`, same parameter / buffer names
(`experts.{i}.gate_up_proj.weight`, `experts.{i}.down_proj.weight`, `gate.weight`,
`gate.e_score_correction_bias`, `shared_experts.*`).  The decoder layer swaps it in by importing
this `Glm4MoeMoE` instead of the reference's at `decode_layer.py:6`.

Differences that are deliberate: the debug safetensors read of
moe.py:153-157 is gone; expert weights live in two stacked tensors (`[E+1, 2I, H]`, `[E+1, H, I]`,
last index = shared expert) of which the per-expert parameters are views; the combine
accumulates in fp32; the loader matches keys exactly (the reference's `startswith` at
mlp.py:96 lets `experts.1` also match `experts.10`..).  All compute runs in the CUDA library
(`csrc/`, C ABI in `include/glm4_moe_b200.h`); there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
from torch import nn

from . import _lib

_VIEW = {"idx": 0, "w": 1, "expert_offsets": 2, "row_token": 3, "slot_row": 4, "logits": 5,
         "act": 6, "y": 7, "num_tiles": 8, "tiles": 9, "x_rows": 10, "row_weight": 11}
_IMPL = {"auto": _lib.IMPL_AUTO, "simt": _lib.IMPL_SIMT, "umma": _lib.IMPL_UMMA}


class _Weight(nn.Module):
    """Stand-in for the reference's linear layers (linear.py:56-155): only `.weight` [out, in]."""

    def __init__(self, view: torch.Tensor):
        super().__init__()
        self.weight = nn.Parameter(view, requires_grad=False)


class Glm4MoeMLP(nn.Module):
    """One expert: `gate_up_proj.weight [2I, H]` (rows [0,I) gate, [I,2I) up — mlp.py:118) and
    `down_proj.weight [H, I]`, both views into the owning block's stacks (mlp.py:27-38)."""

    def __init__(self, gate_up_view: torch.Tensor, down_view: torch.Tensor, hidden_act: str = "silu"):
        super().__init__()
        if hidden_act.lower() != "silu":      # mlp.py:41-44
            raise ValueError(f"Unsupported activation: {hidden_act}. Only silu is supported for now.")
        self.gate_up_proj = _Weight(gate_up_view)
        self.down_proj = _Weight(down_view)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """mlp.py:49-54 for one expert: the tensor-core dense path (`glm4moe_mlp_forward`) for bf16 / fp16
        weights whose shape it covers, else the CUDA-core expert kernels (exact fp32-FMA path)."""
        wgu, wd = self.gate_up_proj.weight, self.down_proj.weight
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        _lib.require_cuda(x2, "x")
        return dense_mlp_forward(x2, wgu, wd, self)


def dense_mlp_forward(x2: torch.Tensor, wgu: torch.Tensor, wd: torch.Tensor, owner=None, impl: str = "auto"):
    """`down(silu(gate(x)) * up(x))` of ONE weight pair (mlp.py:49-54) on `x2 [n, H]`.  `impl`: "auto" picks
    the tcgen05 dense path where `glm4moe_mlp_forward` supports the shape (2-byte dtype, H % 128 == 0,
    I % 8 == 0), "simt" forces the CUDA-core kernels, "umma" insists on the tensor cores."""
    lib = _lib.load()
    n, H = x2.shape
    I = wd.shape[1]
    dt = _lib.DTYPE_CODE[x2.dtype]
    dev = x2.device
    act = torch.empty((n, I), dtype=x2.dtype, device=dev)
    y = torch.empty((n, H), dtype=x2.dtype, device=dev)
    if n == 0:
        return y
    tc_ok = x2.dtype in (torch.bfloat16, torch.float16) and H % 128 == 0 and I % 8 == 0 and wgu.dtype == x2.dtype
    if impl == "umma" and not tc_ok:
        raise _lib.Glm4MoeLibraryError(f"dense tensor-core MLP does not cover H={H}, I={I}, {x2.dtype}")
    st = _lib.stream_ptr(dev)
    if tc_ok and impl != "simt":
        # 4 self re-arming counters, zeroed once per (owner, device)
        sched = getattr(owner, "_dense_sched", None) if owner is not None else None
        if sched is None or sched.device != dev:
            sched = torch.zeros(4, dtype=torch.int32, device=dev)
            if owner is not None:
                object.__setattr__(owner, "_dense_sched", sched)
        with torch.cuda.device(dev):
            _lib.check(lib.glm4moe_mlp_forward(x2.data_ptr(), wgu.data_ptr(), wd.data_ptr(), act.data_ptr(),
                                               y.data_ptr(), n, H, I, dt, sched.data_ptr(), st), "mlp_forward")
        return y
    ntile = (n + 7) // 8
    tiles = torch.tensor([[0, 8 * i, min(8, n - 8 * i), 0] for i in range(ntile)], dtype=torch.int32,
                         device=dev).reshape(-1, 4)
    num = torch.tensor([ntile], dtype=torch.int32, device=dev)
    ones = torch.ones(n, dtype=torch.float32, device=dev)
    _lib.check(lib.glm4moe_expert_gate_up_simt(x2.data_ptr(), wgu.data_ptr(), tiles.data_ptr(),
                                               num.data_ptr(), act.data_ptr(), H, I, 8, dt, st), "gate_up")
    # T*k = n with k = 1 so every row takes its weight (1.0) from `ones`
    _lib.check(lib.glm4moe_expert_down_simt(act.data_ptr(), wd.data_ptr(), tiles.data_ptr(),
                                            num.data_ptr(), ones.data_ptr(), y.data_ptr(), H, I, n, 1,
                                            8, dt, st), "down")
    return y


class Glm4MoeDenseMLP(nn.Module):
    """The DENSE MLP of the decoder stack's first layers (`decode_layer.py:31-36` constructs the reference's
    `Glm4MoeMLP(hidden_size, intermediate_size, hidden_act, prefix=...)`, `mlp.py:12-54`) through the same
    tensor-core grouped GEMM as the experts ("the E = 1 case").  Same constructor, parameter
    names (`gate_up_proj.weight [2I/tp, H]`, `down_proj.weight [H, I/tp]`) and `load_weights(path, prefix)`
    as the reference class; tensor parallel like its `MergedColumnParallelLinear` / `RowParallelLinear`
    (`linear.py:79-155`): rank r holds rows [r*I/tp, (r+1)*I/tp) of gate and of up and the matching columns of
    down, and the partial outputs are summed with `dist.all_reduce` (`linear.py:153-154`)."""

    def __init__(self, hidden_size: int, intermediate_size: int, hidden_act: str = "silu", quant_config=None,
                 reduce_results: bool = True, prefix: str = "", tp_rank: int | None = None, tp_size: int | None = None):
        super().__init__()
        if hidden_act.lower() != "silu":      # mlp.py:41-44
            raise ValueError(f"Unsupported activation: {hidden_act}. Only silu is supported for now.")
        import torch.distributed as dist
        live = dist.is_available() and dist.is_initialized()
        self.tp_size = tp_size if tp_size is not None else (dist.get_world_size() if live else 1)
        self.tp_rank = tp_rank if tp_rank is not None else (dist.get_rank() if live else 0)
        if intermediate_size % self.tp_size:
            raise ValueError(f"intermediate_size={intermediate_size} not divisible by tp_size={self.tp_size}")
        self.hidden_size, self.intermediate_size = hidden_size, intermediate_size
        self.local_intermediate = intermediate_size // self.tp_size
        self.reduce_results = reduce_results
        self.prefix = prefix
        self.impl = "auto"
        self.gate_up_proj = _Weight(torch.empty((2 * self.local_intermediate, hidden_size)))
        self.down_proj = _Weight(torch.empty((hidden_size, self.local_intermediate)))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        _lib.require_cuda(x2, "x")
        y = dense_mlp_forward(x2, self.gate_up_proj.weight, self.down_proj.weight, self, self.impl)
        if self.tp_size > 1 and self.reduce_results:
            import torch.distributed as dist
            dist.all_reduce(y)
        return y.reshape(*x.shape[:-1], self.hidden_size)

    def load_weights(self, model_path: str, prefix: str | None = None):
        """`{prefix}.gate_proj|up_proj|down_proj.weight` of an HF-layout checkpoint (a `.safetensors` file or a
        directory of them), the keys the reference loader reads (`mlp.py:83-131`); one pass, exact-key match,
        this rank's shard only."""
        import os
        from safetensors import safe_open
        prefix = self.prefix if prefix is None else prefix
        files = [model_path] if os.path.isfile(model_path) else \
            [os.path.join(model_path, f) for f in sorted(os.listdir(model_path)) if f.endswith(".safetensors")]
        Il, r = self.local_intermediate, self.tp_rank
        seen = set()
        with torch.no_grad():
            for fn in files:
                with safe_open(fn, framework="pt", device="cpu") as f:
                    keys = set(f.keys())
                    for proj in ("gate_proj", "up_proj", "down_proj"):
                        key = f"{prefix}.{proj}.weight"
                        if key not in keys:
                            continue
                        if proj == "down_proj":
                            t = f.get_slice(key)[:, r * Il:(r + 1) * Il]
                            self.down_proj.weight.copy_(t)
                        else:
                            t = f.get_slice(key)[r * Il:(r + 1) * Il, :]
                            dst = self.gate_up_proj.weight[:Il] if proj == "gate_proj" else self.gate_up_proj.weight[Il:]
                            dst.copy_(t)
                        seen.add(proj)
        missing = {"gate_proj", "up_proj", "down_proj"} - seen
        if missing:
            raise ValueError(f"No tensors {sorted(missing)} with prefix '{prefix}' found in '{model_path}'")


class GlmMoeSelectTopk(nn.Module):
    """Router (moe.py:13-50): returns `(topk_indices int64 [T,k], topk_weights model-dtype [T,k])`.
    Slots come back sorted by ascending expert id (the reference's order is unspecified,
    `torch.topk(sorted=False)`); ties are broken towards the lower expert id."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.top_k = config.num_experts_per_tok
        self.n_routed_experts = config.n_routed_experts
        assert config.n_group == 1          # moe.py:21
        assert config.topk_group == 1       # moe.py:22
        self.routed_scaling_factor = config.routed_scaling_factor
        self.weight = nn.Parameter(torch.empty((self.n_routed_experts, config.hidden_size)),
                                   requires_grad=False)
        self.register_buffer("e_score_correction_bias",
                             torch.zeros((self.n_routed_experts,), dtype=torch.float32))

    def _apply(self, fn, recurse=True):
        """`.half()` / `.bfloat16()` / `.to(dtype)` cast floating-point buffers too (nn.Module._apply);
        the kernels read `e_score_correction_bias` as fp32 (moe.py:26,38), so the buffer follows device
        moves only and keeps its exact fp32 values."""
        bias = self.e_score_correction_bias
        out = super()._apply(fn, recurse)
        moved = self.e_score_correction_bias
        if moved.dtype != torch.float32:
            self.e_score_correction_bias = bias.to(device=moved.device, dtype=torch.float32)
        return out

    def bias_ptr(self) -> int:
        """Device pointer of the fp32 bias, validated (the library reads E * 4 bytes from it)."""
        b = self.e_score_correction_bias
        if b.dtype != torch.float32 or not b.is_contiguous() or b.numel() != self.n_routed_experts:
            raise _lib.Glm4MoeLibraryError(
                f"gate.e_score_correction_bias must be a contiguous fp32 [{self.n_routed_experts}] buffer, "
                f"got {b.dtype} {tuple(b.shape)}")
        _lib.require_cuda(b, "gate.e_score_correction_bias")
        return b.data_ptr()

    def forward(self, hidden_states: torch.Tensor, return_logits: bool = False):
        lib = _lib.load()
        H = self.config.hidden_size
        h = hidden_states.reshape(-1, H).contiguous()
        _lib.require_cuda(h, "hidden_states")
        if h.dtype != self.weight.dtype:
            raise _lib.Glm4MoeLibraryError(f"hidden_states dtype {h.dtype} != router dtype {self.weight.dtype}")
        T, E, k = h.shape[0], self.n_routed_experts, self.top_k
        dt = _lib.DTYPE_CODE[h.dtype]
        logits = torch.empty((T, E), dtype=h.dtype, device=h.device)
        idx = torch.empty((T, k), dtype=torch.int32, device=h.device)
        w = torch.empty((T, k), dtype=torch.float32, device=h.device)
        st = _lib.stream_ptr(h.device)
        _lib.check(lib.glm4moe_router_logits(h.data_ptr(), self.weight.data_ptr(), logits.data_ptr(), T, H, E, dt, st),
                   "router_logits")
        _lib.check(lib.glm4moe_router_select(logits.data_ptr(), self.bias_ptr(), T, E, k,
                                             float(self.routed_scaling_factor), dt, idx.data_ptr(), w.data_ptr(), st),
                   "router_select")
        out = (idx.to(torch.int64), w.to(h.dtype))
        return out + (logits,) if return_logits else out


class _ForeignCudaBuffer:
    """uint8 view of device memory the CUDA library allocated (zero-copy via __cuda_array_interface__)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class Glm4MoeMoE(nn.Module):
    """GLM4-MoE sparse block (moe.py:52-176).

    Extra keyword arguments (all default to the reference's single-GPU behaviour):
      max_tokens   initial workspace capacity in tokens (grows on demand outside graph capture)
      ep_rank/ep_size  expert-parallel shard: this module then owns experts
                   [rank*E/size, (rank+1)*E/size) and `forward` returns the *partial* sum over the
                   local experts (+ the shared expert on this rank's 1/size slice of the tokens,
                   `with_shared = 2`); see distributed.py
      impl         "auto" | "simt" | "umma": expert-GEMM implementation

    Class-level construction defaults, for hosts that call `Glm4MoeMoE(config)` with no room for keyword
    arguments (the reference decoder layer, decode_layer.py:29):
      default_weight_format  "native" | "fp8": "fp8" allocates the e4m3 stacks + block scales directly
                   and no model-dtype stacks (a 46-layer GLM-4.5-Air holds 45 of these blocks: 199 GB in
                   bf16, 101 GB in e4m3)
      share_workspace  True: blocks of one device with the same shape share ONE workspace allocation
                   (layers of a model run one after the other on one stream; every in-kernel counter
                   re-arms itself).  A private workspace per layer costs 2.8 GB at 16 k prefill tokens.
      default_ep   None | (ep_rank, ep_size, peer_workspace): expert-parallel shard of a block constructed
                   without keyword arguments (engine_shim.py sets it per process once torch.distributed is up)
      default_max_tokens  initial workspace capacity of such a block
    """

    default_weight_format = "native"
    share_workspace = False
    default_ep = None
    default_max_tokens = None
    _workspace_pool: dict = {}
    # share_workspace + peer_workspace: ONE cudaMalloc'd, IPC-exported workspace per (device, layout), used by every
    # expert-parallel block of the process.  key -> {"ptr", "handle", "tensor", "users" (WeakSet), "used" (a forward ran)}
    _peer_workspace_pool: dict = {}

    def __init__(self, config, *, max_tokens: int | None = None, ep_rank: int | None = None,
                 ep_size: int | None = None, impl: str = "auto", peer_workspace: bool | None = None):
        super().__init__()
        self.config = config
        cls = type(self)
        if ep_rank is None and ep_size is None and cls.default_ep is not None:
            ep_rank, ep_size = cls.default_ep[0], cls.default_ep[1]
            if peer_workspace is None:
                peer_workspace = cls.default_ep[2]
        ep_rank, ep_size = ep_rank or 0, ep_size or 1
        if max_tokens is None:
            max_tokens = cls.default_max_tokens or 256
        E, H, I = config.n_routed_experts, config.hidden_size, config.moe_intermediate_size
        if E % ep_size != 0:
            raise ValueError(f"n_routed_experts={E} not divisible by ep_size={ep_size}")
        self.ep_rank, self.ep_size = ep_rank, ep_size
        self.expert_begin = ep_rank * (E // ep_size)
        self.expert_end = self.expert_begin + E // ep_size
        # single GPU: shared expert on every token; expert-parallel: on this rank's 1/R token slice
        self.with_shared = 1 if ep_size == 1 else 2
        self.impl = impl
        # peer_workspace: the layer workspace is a cudaMalloc'd, IPC-exportable region (fixed capacity)
        # that the other expert-parallel ranks map and read (distributed.py layout "p2p")
        self.peer_workspace = bool(peer_workspace)
        self._ipc_ptr = None
        self._ipc_handle = None
        self._peer_pool_entry = None
        n_local = self.expert_end - self.expert_begin
        # stacked storage, created under the process-wide default dtype/device like the
        # reference's parameters (model_runner.py:29-31)
        fp8_only = type(self).default_weight_format in ("fp8", "fp8_pow2")
        if fp8_only:
            if I % 128 or H % 128:
                raise ValueError(f"block-128 FP8 needs dimensions divisible by 128, got I={I}, H={H}")
            self._w_gate_up = torch.empty((0,))
            self._w_down = torch.empty((0,))
            gu = [self._w_gate_up] * (n_local + 1)
            dn = [self._w_down] * (n_local + 1)
        else:
            self._w_gate_up = torch.empty((n_local + 1, 2 * I, H))
            self._w_down = torch.empty((n_local + 1, H, I))
            gu, dn = self._w_gate_up, self._w_down
        self.experts = nn.ModuleList([Glm4MoeMLP(gu[i], dn[i], config.hidden_act) for i in range(n_local)])
        self.gate = GlmMoeSelectTopk(config)
        self.shared_experts = Glm4MoeMLP(gu[n_local], dn[n_local], config.hidden_act)
        self.weight_format = "native"
        # FP8 tier with power-of-two (UE8M0) block scales: activations get power-of-two scales too and the
        # prefill GEMMs use the hardware block-scaled MMA (csrc/umma_fp8_mx.cu)
        self.fp8_pow2_scales = type(self).default_weight_format == "fp8_pow2"
        self._wq_gate_up = self._wq_down = self._s_gate_up = self._s_down = None
        if fp8_only:
            f8 = torch.float8_e4m3fn
            self._wq_gate_up = torch.empty((n_local + 1, 2 * I, H), dtype=f8)
            self._wq_down = torch.empty((n_local + 1, H, I), dtype=f8)
            self._s_gate_up = torch.empty((n_local + 1, 2 * I // 128, H // 128), dtype=torch.float32)
            self._s_down = torch.empty((n_local + 1, H // 128, I // 128), dtype=torch.float32)
            self.weight_format = "fp8"
        self._max_tokens = int(max_tokens)
        self._handle = None
        self._workspace = None
        self._host = None

    # ---- storage management -------------------------------------------------------------
    def _mlps(self):
        return list(self.experts) + [self.shared_experts]

    def _apply(self, fn, recurse=True):
        """Keep per-expert parameters aliased to the stacks across .to()/.cuda()/.half().  Only the two
        stacks are moved / cast: the per-expert views are parked on an empty tensor while
        `nn.Module._apply` walks the parameters (it would otherwise copy every view on its own, a second
        full copy of the stack on the target device) and re-aliased afterwards."""
        self._destroy_handle()
        self._w_gate_up = fn(self._w_gate_up)
        self._w_down = fn(self._w_down)
        if self._wq_gate_up is not None:      # fp8 stacks follow device moves only (never re-cast)
            dev = self._w_gate_up.device
            self._wq_gate_up, self._wq_down = self._wq_gate_up.to(dev), self._wq_down.to(dev)
            self._s_gate_up, self._s_down = self._s_gate_up.to(dev), self._s_down.to(dev)
        park = torch.empty((0,), dtype=self._w_gate_up.dtype, device=self._w_gate_up.device)
        for mlp in self._mlps():
            mlp.gate_up_proj.weight.data = park
            mlp.down_proj.weight.data = park
        out = super()._apply(fn, recurse)
        self._alias_experts()
        return out

    def _alias_experts(self):
        """Point `experts.{i}.*.weight` / `shared_experts.*.weight` at their slices of the stacks (or at
        the empty stack after `quantize_fp8(drop_native=True)`)."""
        dropped = self._w_gate_up.dim() != 3
        for i, mlp in enumerate(self._mlps()):
            mlp.gate_up_proj.weight.data = self._w_gate_up if dropped else self._w_gate_up[i]
            mlp.down_proj.weight.data = self._w_down if dropped else self._w_down[i]

    def _destroy_handle(self):
        if getattr(self, "_handle", None) is not None:
            _lib.load().glm4moe_layer_destroy(self._handle)
        self._handle = None
        for _, users in type(self)._workspace_pool.values():
            users.discard(self)
        self._workspace = None
        self._host = None
        ent = getattr(self, "_peer_pool_entry", None)
        if ent is not None:
            # shared peer workspace: freed with its last user (peers unmap first: ExpertParallelMoE.close_p2p)
            ent["users"].discard(self)
            self._peer_pool_entry = None
            if not len(ent["users"]):
                pool = type(self)._peer_workspace_pool
                for k_ in [k_ for k_, v in pool.items() if v is ent]:
                    del pool[k_]
                ent["tensor"] = None
                _lib.load().glm4moe_p2p_free(ent["ptr"])
        elif getattr(self, "_ipc_ptr", None):
            _lib.load().glm4moe_p2p_free(self._ipc_ptr)
        self._ipc_ptr = None
        self._ipc_handle = None

    def __del__(self):
        try:
            self._destroy_handle()
        except Exception:
            pass

    def _lib_config(self, max_tokens: int) -> _lib.Config:
        c = self.config
        fmt = _lib.W_NATIVE
        if self.weight_format == "fp8":
            fmt = _lib.W_FP8_BLOCK128_POW2 if self.fp8_pow2_scales else _lib.W_FP8_BLOCK128
        return _lib.Config(c.hidden_size, c.n_routed_experts, c.num_experts_per_tok, c.moe_intermediate_size,
                           float(c.routed_scaling_factor), _lib.DTYPE_CODE[self.gate.weight.dtype], fmt,
                           max_tokens, self.expert_begin, self.expert_end, int(self.with_shared),
                           _IMPL[self.impl])

    def _ensure_layer(self, T: int):
        if self._handle is not None and T <= self._max_tokens:
            return
        lib = _lib.load()
        _lib.require_cuda(self.gate.weight, "weights")
        if torch.cuda.is_current_stream_capturing():
            raise _lib.Glm4MoeLibraryError(
                f"workspace for {T} tokens must exist before CUDA-graph capture (run a warm-up forward first)")
        if self.peer_workspace and self._handle is not None:
            raise _lib.Glm4MoeLibraryError(
                f"peer-mapped workspace holds {self._max_tokens} tokens, got {T}: it is mapped by the other "
                "ranks and cannot grow; construct the block with a larger max_tokens")
        self._destroy_handle()
        cap = max(self._max_tokens, 1)
        while cap < T:
            cap *= 2
        self._max_tokens = cap
        cfg = self._lib_config(cap)
        nbytes = lib.glm4moe_workspace_bytes(C.byref(cfg))
        if nbytes == 0:
            raise _lib.Glm4MoeLibraryError("unsupported GLM4-MoE configuration for the CUDA library")
        dev = self.gate.weight.device
        if self.peer_workspace and type(self).share_workspace:
            # One peer-mapped workspace for all the expert-parallel blocks of this process (a 46-layer model would
            # otherwise hold 45 of them: 3 GB each at 16 k tokens).  The layers of a model run one after the other
            # on one stream and every cross-GPU counter (epoch, parity, ready / done flags) lives IN the workspace,
            # so to the protocol of ep_combine.cu the next layer is simply the next call.  `glm4moe_layer_create`
            # zeroes those counters: every sharer must be created before the first forward on the workspace.
            key = (str(dev), nbytes, tuple(getattr(cfg, f[0]) for f in cfg._fields_ if f[0] not in ("impl",)))
            pool = type(self)._peer_workspace_pool
            ent = pool.get(key)
            if ent is None:
                import weakref
                ptr, hbuf = C.c_void_p(), C.create_string_buffer(64)
                with torch.cuda.device(dev):
                    _lib.check(lib.glm4moe_p2p_alloc(nbytes + 1024, C.byref(ptr), hbuf), "p2p_alloc")
                ent = pool[key] = {"ptr": ptr.value, "handle": hbuf.raw, "users": weakref.WeakSet(), "used": False,
                                   "tensor": torch.as_tensor(_ForeignCudaBuffer(ptr.value, nbytes + 1024), device=dev)}
            elif ent["used"]:
                raise _lib.Glm4MoeLibraryError(
                    "a shared peer-mapped workspace already ran a forward: creating another block on it would reset "
                    "its cross-GPU counters. Call setup_p2p() on every expert-parallel block before the first forward")
            ent["users"].add(self)
            self._peer_pool_entry = ent
            self._ipc_ptr, self._ipc_handle, self._workspace = ent["ptr"], ent["handle"], ent["tensor"]
        elif self.peer_workspace:
            ptr, hbuf = C.c_void_p(), C.create_string_buffer(64)
            with torch.cuda.device(dev):
                _lib.check(lib.glm4moe_p2p_alloc(nbytes + 1024, C.byref(ptr), hbuf), "p2p_alloc")
            self._ipc_ptr, self._ipc_handle = ptr.value, hbuf.raw
            self._workspace = torch.as_tensor(_ForeignCudaBuffer(ptr.value, nbytes + 1024), device=dev)
        elif type(self).share_workspace:
            # one allocation per (device, layout): the layout is a function of the config alone, so every
            # block with this key sees the same counters at the same offsets
            key = (str(dev), nbytes, tuple(getattr(cfg, f[0]) for f in cfg._fields_ if f[0] not in ("impl",)))
            pool = type(self)._workspace_pool
            for k_ in [k_ for k_, (_, users) in pool.items() if k_[0] == str(dev) and not len(users)]:
                del pool[k_]                       # grown out of: nobody holds it any more
            if key not in pool:
                import weakref
                pool[key] = (torch.empty(nbytes + 1024, dtype=torch.uint8, device=dev), weakref.WeakSet())
            self._workspace = pool[key][0]
            pool[key][1].add(self)
        else:
            self._workspace = torch.empty(nbytes + 1024, dtype=torch.uint8, device=dev)
        base = (self._workspace.data_ptr() + 1023) // 1024 * 1024
        handle = C.c_void_p()
        if self.weight_format == "fp8":
            wgu, wd = self._wq_gate_up.data_ptr(), self._wq_down.data_ptr()
            sgu, sd = self._s_gate_up.data_ptr(), self._s_down.data_ptr()
        else:
            wgu, wd, sgu, sd = self._w_gate_up.data_ptr(), self._w_down.data_ptr(), None, None
        with torch.cuda.device(dev):
            _lib.check(lib.glm4moe_layer_create(C.byref(cfg), self.gate.weight.data_ptr(),
                                                self.gate.bias_ptr(), wgu, wd, sgu, sd,
                                                base, nbytes, C.byref(handle)), "layer_create")
        self._handle = handle
        self._cfg = cfg

    def quantize_fp8(self, drop_native: bool = False, pow2_scales: bool = False):
        """Convert the expert stacks to e4m3 + 128x128 fp32 block scales (fp8.py) and switch the
        block to the FP8 tensor-core path.  Router weights and activations keep the model dtype.
        `pow2_scales`: power-of-two (UE8M0) scales for weights and activations, which the prefill kernel
        applies in hardware (tcgen05 block-scaled MMA)."""
        from . import fp8
        self._destroy_handle()
        self._wq_gate_up, self._s_gate_up = fp8.quantize_block128(self._w_gate_up, pow2_scales)
        self._wq_down, self._s_down = fp8.quantize_block128(self._w_down, pow2_scales)
        self.weight_format = "fp8"
        self.fp8_pow2_scales = bool(pow2_scales)
        if drop_native:
            dev, dt = self._w_gate_up.device, self._w_gate_up.dtype
            self._w_gate_up = torch.empty((0,), device=dev, dtype=dt)
            self._w_down = torch.empty((0,), device=dev, dtype=dt)
            self._alias_experts()
        return self

    def dequantized_weights(self) -> dict:
        """fp32 stacks the FP8 path is equivalent to (oracle input for the quant-tier parity test)."""
        from . import fp8
        return {"w_gate_up": fp8.dequantize_block128(self._wq_gate_up, self._s_gate_up),
                "w_down": fp8.dequantize_block128(self._wq_down, self._s_down)}

    # ---- reference API ----------------------------------------------------------------------
    def forward(self, hidden_states: torch.Tensor, shared_range: tuple[int, int] | None = None) -> torch.Tensor:
        """moe.py:128-176: returns a fresh 2-D `[T, H]` tensor; the input is not modified.
        `shared_range=(lo, hi)` restricts the shared expert to those tokens (expert-parallel dispatch)."""
        h = hidden_states.reshape(-1, hidden_states.shape[-1])
        if not h.is_contiguous():
            h = h.contiguous()
        _lib.require_cuda(h, "hidden_states")
        if h.dtype != self.gate.weight.dtype:
            raise _lib.Glm4MoeLibraryError(f"hidden_states dtype {h.dtype} != model dtype {self.gate.weight.dtype}")
        T = h.shape[0]
        out = torch.empty_like(h)
        if T == 0:
            return out
        self._ensure_layer(T)
        with torch.cuda.device(h.device), _lib.nvtx_range("glm4moe.forward", T):
            if shared_range is None:
                _lib.check(_lib.load().glm4moe_layer_forward(self._handle, h.data_ptr(), out.data_ptr(), T,
                                                            _lib.stream_ptr(h.device)), "layer_forward")
            else:
                _lib.check(_lib.load().glm4moe_layer_forward_ex(self._handle, h.data_ptr(), out.data_ptr(), T,
                                                               int(shared_range[0]), int(shared_range[1]),
                                                               _lib.stream_ptr(h.device)), "layer_forward_ex")
        return out

    def forward_prenorm(self, hidden_states: torch.Tensor, residual: torch.Tensor | None, norm_weight: torch.Tensor,
                        eps: float):
        """`decode_layer.py:98-99` as one library call (`glm4moe_layer_forward_prenorm`):
        `hidden, residual = post_attention_layernorm(hidden, residual); hidden = mlp(hidden)`.
        Returns `(out [T, H], residual_out)`; `residual=None` is the plain RMSNorm form (`layernorm.py:17-26`)
        and returns `(out, None)`.  The normalised tokens stay in the layer workspace; results are bit-identical
        to `RMSNorm` + `forward`."""
        h, r, T = self._prenorm_args(hidden_states, residual, norm_weight)
        out = torch.empty_like(h)
        res_out = torch.empty_like(h) if r is not None else None
        if T == 0:
            return out, res_out
        self._ensure_layer(T)
        with torch.cuda.device(h.device), _lib.nvtx_range("glm4moe.forward_prenorm", T):
            _lib.check(_lib.load().glm4moe_layer_forward_prenorm(
                self._handle, h.data_ptr(), _lib.ptr(r), norm_weight.data_ptr(), float(eps),
                _lib.ptr(res_out), out.data_ptr(), T, _lib.stream_ptr(h.device)), "layer_forward_prenorm")
        return out, res_out

    def _prenorm_args(self, hidden_states, residual, norm_weight):
        h = hidden_states.reshape(-1, hidden_states.shape[-1])
        if not h.is_contiguous():
            h = h.contiguous()
        _lib.require_cuda(h, "hidden_states")
        md = self.gate.weight.dtype
        if h.dtype != md or norm_weight.dtype != md:
            raise _lib.Glm4MoeLibraryError(f"hidden_states {h.dtype} / norm weight {norm_weight.dtype} != model dtype {md}")
        if norm_weight.numel() != h.shape[-1] or not norm_weight.is_contiguous():
            raise _lib.Glm4MoeLibraryError("norm weight must be a contiguous [H] tensor")
        r = None
        if residual is not None:
            r = residual.reshape(-1, residual.shape[-1])
            if not r.is_contiguous():
                r = r.contiguous()
            if r.shape != h.shape or r.dtype != h.dtype:
                raise _lib.Glm4MoeLibraryError("residual must match hidden_states in shape and dtype")
        return h, r, h.shape[0]

    STAGES = ("router_logits", "router_select", "permute", "gather", "gate_up", "down", "combine", "total")

    def forward_timed(self, hidden_states: torch.Tensor):
        """`forward` with a CUDA event between stages; returns (out, {stage: ms})."""
        h = hidden_states.reshape(-1, hidden_states.shape[-1]).contiguous()
        _lib.require_cuda(h, "hidden_states")
        T = h.shape[0]
        out = torch.empty_like(h)
        self._ensure_layer(T)
        ms = (C.c_float * 8)()
        with torch.cuda.device(h.device):
            _lib.check(_lib.load().glm4moe_layer_forward_timed(self._handle, h.data_ptr(), out.data_ptr(), T,
                                                              _lib.stream_ptr(h.device), ms), "layer_forward_timed")
        return out, dict(zip(self.STAGES, list(ms)))

    def forward_host(self, hidden_states_host: torch.Tensor, out_host: torch.Tensor | None = None) -> torch.Tensor:
        """Host-buffer entry (`glm4moe_layer_forward_host`): H2D copy, forward, D2H copy, sync."""
        h = hidden_states_host.reshape(-1, hidden_states_host.shape[-1])
        if h.is_cuda or not h.is_contiguous():
            raise _lib.Glm4MoeLibraryError("forward_host expects a contiguous host tensor")
        T, H = h.shape
        self._ensure_layer(T)
        dev = self.gate.weight.device
        if self._host is None or self._host[0].shape[0] < T:
            self._host = (torch.empty((self._max_tokens, H), dtype=h.dtype, device=dev),
                          torch.empty((self._max_tokens, H), dtype=h.dtype, device=dev))
        if out_host is None:
            out_host = torch.empty((T, H), dtype=h.dtype, pin_memory=True)
        with torch.cuda.device(dev):
            _lib.check(_lib.load().glm4moe_layer_forward_host(self._handle, h.data_ptr(), out_host.data_ptr(), T,
                                                             self._host[0].data_ptr(), self._host[1].data_ptr(),
                                                             _lib.stream_ptr(dev)), "layer_forward_host")
        return out_host

    def load_weights(self, path: str, prefix: str):
        """moe.py:71-126 + mlp.py:83-131: HF keys -> parameters, one pass, exact key match.

        Keys: `{p}.gate.weight`, `{p}.gate.e_score_correction_bias`,
        `{p}.experts.{i}.{gate_proj,up_proj,down_proj}.weight`, `{p}.shared_experts.*`;
        gate_proj and up_proj are concatenated along dim 0 (mlp.py:118), nothing is transposed.
        An expert-parallel shard reads only its own experts.  FP8 checkpoints (e4m3 `weight` +
        fp32 `weight_scale_inv [rows/128, cols/128]`, the block format of fp8.py) fill the
        quantised stacks directly and switch the block to the e4m3 kernels.
        """
        from safetensors import safe_open
        if os.path.isdir(path):
            files = sorted(os.path.join(path, f) for f in os.listdir(path) if f.endswith(".safetensors"))
        else:
            files = [path]
        I, H = self.config.moe_intermediate_size, self.config.hidden_size
        n_local = self.expert_end - self.expert_begin
        found = 0
        seen, seen_scale = set(), set()
        f8 = torch.float8_e4m3fn

        def fp8_stacks():
            if self._wq_gate_up is None:
                if I % 128 or H % 128:
                    raise ValueError(f"block-128 FP8 needs dimensions divisible by 128, got I={I}, H={H}")
                dev = self._w_gate_up.device
                self._wq_gate_up = torch.empty((n_local + 1, 2 * I, H), dtype=f8, device=dev)
                self._wq_down = torch.empty((n_local + 1, H, I), dtype=f8, device=dev)
                self._s_gate_up = torch.empty((n_local + 1, 2 * I // 128, H // 128), dtype=torch.float32, device=dev)
                self._s_down = torch.empty((n_local + 1, H // 128, I // 128), dtype=torch.float32, device=dev)

        def put(dst, t, key):
            assert t.shape == dst.shape, f"{key}: {tuple(t.shape)} != {tuple(dst.shape)}"
            if t.dtype == f8:
                dst.view(torch.uint8).copy_(t.view(torch.uint8))     # bytes, no conversion
            else:
                dst.copy_(t)

        with torch.no_grad():
            for fn in files:
                with safe_open(fn, framework="pt", device="cpu") as f:
                    for key in f.keys():
                        if not key.startswith(prefix + "."):
                            continue
                        found += 1
                        rest = key[len(prefix) + 1:]
                        if rest == "gate.weight":
                            self.gate.weight.data.copy_(f.get_tensor(key))
                            continue
                        if rest == "gate.e_score_correction_bias":
                            self.gate.e_score_correction_bias.copy_(f.get_tensor(key))
                            continue
                        parts = rest.split(".")
                        if parts[0] == "experts" and len(parts) == 4:
                            e = int(parts[1])
                            if not (self.expert_begin <= e < self.expert_end):
                                continue
                            slot, proj, leaf = e - self.expert_begin, parts[2], parts[3]
                        elif parts[0] == "shared_experts" and len(parts) == 3:
                            slot, proj, leaf = n_local, parts[1], parts[2]
                        else:
                            continue
                        if proj not in ("gate_proj", "up_proj", "down_proj") or leaf not in ("weight", "weight_scale_inv"):
                            continue
                        t = f.get_tensor(key)
                        if leaf == "weight_scale_inv":
                            fp8_stacks()
                            sI = I // 128
                            dst = {"gate_proj": self._s_gate_up[slot, :sI], "up_proj": self._s_gate_up[slot, sI:],
                                   "down_proj": self._s_down[slot]}[proj]
                            put(dst, t.float(), key)
                            seen_scale.add((slot, proj))
                            continue
                        if t.dtype == f8:
                            fp8_stacks()
                            gu, dn = self._wq_gate_up, self._wq_down
                        else:
                            gu, dn = self._w_gate_up, self._w_down
                        dst = {"gate_proj": gu[slot, :I], "up_proj": gu[slot, I:], "down_proj": dn[slot]}[proj]
                        put(dst, t, key)
                        seen.add((slot, proj, t.dtype == f8))
        if not found:
            raise ValueError(f"No tensors with prefix '{prefix}' found in '{path}'")   # moe.py:94-95
        is_fp8 = any(q for _, _, q in seen)
        for slot in range(n_local + 1):
            for proj in ("gate_proj", "up_proj", "down_proj"):
                assert (slot, proj, is_fp8) in seen, f"{proj}.weight missing for expert slot {slot}"   # mlp.py:113-115
                assert not is_fp8 or (slot, proj) in seen_scale, f"{proj}.weight_scale_inv missing for expert slot {slot}"
        if is_fp8:
            from . import fp8
            self._destroy_handle()
            self.weight_format = "fp8"
            # a checkpoint whose block scales are all powers of two (UE8M0) takes the hardware-scaled path
            self.fp8_pow2_scales = fp8.is_pow2(self._s_gate_up) and fp8.is_pow2(self._s_down)

    # ---- helpers used by tests / bench -------------------------------------------------------
    def init_random_(self, seed: int, std: float = 0.02, bias_high: float = 0.1, chunk: int = 16):
        """Random-init in place (no checkpoint, no network): weights ~ N(0, std^2), bias ~ U(0, bias_high)
        as in synth.py, drawn on the block's own device.  Every expert is drawn from its own generator seed
        (its GLOBAL expert id; the shared expert, the gate and the bias follow), so an expert-parallel shard
        holds exactly the experts of the single-GPU block.  FP8 blocks quantise `chunk` experts at a time
        (fp8.quantize_block128), so no full model-dtype stack exists."""
        from . import fp8
        dev, md = self.gate.weight.device, self.gate.weight.dtype
        gen = torch.Generator(device=dev)
        c = self.config
        H, I, E = c.hidden_size, c.moe_intermediate_size, c.n_routed_experts
        n = self.expert_end - self.expert_begin + 1
        ids = list(range(self.expert_begin, self.expert_end)) + [E]

        def draw(shape, ident):
            gen.manual_seed(seed * 100003 + ident)
            return (torch.randn(shape, generator=gen, device=dev, dtype=torch.float32) * std).to(md)

        with torch.no_grad():
            self._destroy_handle()
            for s0 in range(0, n, chunk):
                m = min(chunk, n - s0)
                gu = torch.stack([draw((2 * I, H), 2 * ids[s0 + j]) for j in range(m)])
                dn = torch.stack([draw((H, I), 2 * ids[s0 + j] + 1) for j in range(m)])
                if self.weight_format == "fp8":
                    q, sc = fp8.quantize_block128(gu, self.fp8_pow2_scales)
                    self._wq_gate_up[s0:s0 + m].view(torch.uint8).copy_(q.view(torch.uint8))
                    self._s_gate_up[s0:s0 + m].copy_(sc)
                    q, sc = fp8.quantize_block128(dn, self.fp8_pow2_scales)
                    self._wq_down[s0:s0 + m].view(torch.uint8).copy_(q.view(torch.uint8))
                    self._s_down[s0:s0 + m].copy_(sc)
                else:
                    self._w_gate_up[s0:s0 + m].copy_(gu)
                    self._w_down[s0:s0 + m].copy_(dn)
            self.gate.weight.data.copy_(draw((E, H), 2 * E + 2))
            gen.manual_seed(seed * 100003 + 2 * E + 3)
            self.gate.e_score_correction_bias.copy_(
                torch.rand(E, generator=gen, device=dev, dtype=torch.float32) * bias_high)
        return self

    def load_stacked(self, weights: dict):
        """Fill from the stacked synthetic dict of `synth.make_weights` (full E+1 stacks)."""
        with torch.no_grad():
            sl = list(range(self.expert_begin, self.expert_end)) + [self.config.n_routed_experts]
            self._w_gate_up.copy_(weights["w_gate_up"][sl])
            self._w_down.copy_(weights["w_down"][sl])
            self.gate.weight.data.copy_(weights["gate_weight"])
            self.gate.e_score_correction_bias.copy_(weights["bias"])

    def load_local_stacks(self, weights: dict):
        """Fill from stacks that already hold only this rank's experts + the shared expert."""
        with torch.no_grad():
            self._w_gate_up.copy_(weights["w_gate_up"])
            self._w_down.copy_(weights["w_down"])
            self.gate.weight.data.copy_(weights["gate_weight"])
            self.gate.e_score_correction_bias.copy_(weights["bias"])

    def workspace_view(self, name: str, T: int | None = None) -> torch.Tensor:
        """Typed copy of a workspace region after a forward (parity triage)."""
        lib = _lib.load()
        p, n = C.c_void_p(), C.c_size_t()
        _lib.check(lib.glm4moe_layer_view(self._handle, _VIEW[name], C.byref(p), C.byref(n)), "layer_view")
        base = (self._workspace.data_ptr() + 1023) // 1024 * 1024
        off = p.value - base
        raw = self._workspace[(base - self._workspace.data_ptr()) + off:][:n.value]
        c = self.config
        k, E, H, I = c.num_experts_per_tok, c.n_routed_experts, c.hidden_size, c.moe_intermediate_size
        md = self.gate.weight.dtype
        if name in ("idx", "row_token", "slot_row", "expert_offsets", "num_tiles", "tiles"):
            t = raw.view(torch.int32)
            if name in ("idx", "slot_row"):
                t = t.view(-1, k)
            if name == "tiles":
                t = t.view(-1, 4)
        elif name in ("w", "row_weight"):
            t = raw.view(torch.float32)
            if name == "w":
                t = t.view(-1, k)
        elif name == "logits":
            t = raw.view(md).view(-1, E)
        elif name == "act":
            t = raw.view(md).view(-1, I)
        else:
            t = raw.view(md).view(-1, H)
        if T is not None and name in ("idx", "w", "slot_row", "logits"):
            t = t[:T]
        return t.clone()

    @property
    def last_launches(self) -> int:
        return _lib.load().glm4moe_layer_last_launches(self._handle) if self._handle else 0

    @property
    def last_impl(self) -> str:
        code = _lib.load().glm4moe_layer_last_impl(self._handle) if self._handle else 0
        return {0: "none", 1: "simt", 2: "umma"}[code]

    @property
    def last_fused(self) -> bool:
        """True when the last forward ran the experts as the single persistent decode kernel."""
        return bool(_lib.load().glm4moe_layer_last_fused(self._handle)) if self._handle else False

    # ---- expert parallelism over peer memory (distributed.py layout "p2p") -------------------------
    def ipc_handle(self, max_tokens: int | None = None) -> tuple[bytes, int]:
        """(64-byte cudaIpc handle of the workspace allocation, offset of the 1 KB-aligned workspace base)."""
        if not self.peer_workspace:
            raise _lib.Glm4MoeLibraryError("construct the block with peer_workspace=True")
        self._ensure_layer(max_tokens or self._max_tokens)
        base = (self._ipc_ptr + 1023) // 1024 * 1024
        return self._ipc_handle, base - self._ipc_ptr

    def workspace_base(self) -> int:
        return (self._workspace.data_ptr() + 1023) // 1024 * 1024

    def forward_ep_p2p(self, hidden_states: torch.Tensor, peer_bases) -> torch.Tensor:
        """Local experts + the peer-memory combine (`glm4moe_layer_forward_ep_p2p`): returns the full
        replicated `[T, H]`, bit-identical to the single-GPU block.  `peer_bases`: ctypes array of the
        workspace base of every rank as mapped in this process."""
        h = hidden_states.reshape(-1, hidden_states.shape[-1])
        if not h.is_contiguous():
            h = h.contiguous()
        _lib.require_cuda(h, "hidden_states")
        if h.dtype != self.gate.weight.dtype:
            raise _lib.Glm4MoeLibraryError(f"hidden_states dtype {h.dtype} != model dtype {self.gate.weight.dtype}")
        T = h.shape[0]
        out = torch.empty_like(h)
        if T == 0:
            return out
        self._ensure_layer(T)
        if self._peer_pool_entry is not None:
            self._peer_pool_entry["used"] = True
        with torch.cuda.device(h.device), _lib.nvtx_range("glm4moe.forward_ep_p2p", T):
            _lib.check(_lib.load().glm4moe_layer_forward_ep_p2p(self._handle, h.data_ptr(), out.data_ptr(), T,
                                                               peer_bases, self.ep_rank, self.ep_size,
                                                               _lib.stream_ptr(h.device)), "layer_forward_ep_p2p")
        return out

    def forward_ep_p2p_prenorm(self, partial: torch.Tensor, residual: torch.Tensor | None, norm_weight: torch.Tensor,
                               eps: float, peer_bases, comm):
        """The tensor-parallel boundary + norm + expert-parallel block in one library call
        (`glm4moe_layer_forward_ep_p2p_prenorm`): `partial` is this rank's o_proj output BEFORE the
        all-reduce of `linear.py:153-154`; `comm` is the process's `distributed.PeerComm`.  Returns the
        replicated `(out, residual_out)`."""
        h, r, T = self._prenorm_args(partial, residual, norm_weight)
        out = torch.empty_like(h)
        res_out = torch.empty_like(h) if r is not None else None
        if T == 0:
            return out, res_out
        self._ensure_layer(T)
        if self._peer_pool_entry is not None:
            self._peer_pool_entry["used"] = True
        with torch.cuda.device(h.device), _lib.nvtx_range("glm4moe.forward_ep_p2p_prenorm", T):
            _lib.check(_lib.load().glm4moe_layer_forward_ep_p2p_prenorm(
                self._handle, h.data_ptr(), _lib.ptr(r), norm_weight.data_ptr(), float(eps), _lib.ptr(res_out),
                out.data_ptr(), T, peer_bases, comm.bases, comm.max_tokens, self.ep_rank, self.ep_size,
                _lib.stream_ptr(h.device)), "layer_forward_ep_p2p_prenorm")
        return out, res_out
