"""Expert-parallel GLM4-MoE block over the GPUs of one box (one process per GPU, NCCL/NVLink).

The reference has no expert parallelism: its only distributed hook on this path is the dead
`dist.all_reduce` in `RowParallelLinear.forward` (`nanovllm/layers/linear.py:153-154`, `tp_size`
is pinned to 1 by `nanovllm/distributed/parallel_state.py:1-9`).  This module is the layer
north_star puts in `nanovllm/distributed`: experts are sharded contiguously, rank r owns
`[r*E/R, (r+1)*E/R)`; the router weights are replicated and routing is recomputed identically
on every rank (deterministic kernels), so no routing metadata is exchanged.

Three layouts:

* ``p2p`` (layout A over peer memory, the default on GPUs of one box) — tokens replicated like
  ``allreduce``; every rank's layer workspace is a cudaIpc-shared region, and after the local experts
  ONE kernel per rank reads, for every token, its k weighted expert rows (+ the shared-expert row)
  from the ranks that own them over NVLink and sums them in ascending expert order in fp32
  (`csrc/ep_combine.cu`).  No NCCL call on the data path, CUDA-graph capturable, and the result is
  bit-identical to the single-GPU block.  The handles are exchanged once, on the host, with
  `all_gather_object`.

* ``allreduce`` (layout A) — activations entering the block are replicated (TP attention ends
  in an all-reduce).  Each rank runs its local experts over all T tokens plus the shared expert
  on its 1/R token slice, producing a partial `[T, H]`; one all-reduce sums the partials.  No
  dispatch, no host sync, CUDA-graph capturable.
* ``alltoall`` (layout B) — rank r owns tokens `[r*T/R, (r+1)*T/R)`, sends each owned token once
  to every rank hosting one of its experts (all-to-all-v), the hosts return the weighted partial
  sums over their local experts (reverse all-to-all-v), the owner adds them (and its shared
  expert) and an all-gather rebuilds the replicated `[T, H]`.  Split sizes are read on the host
  (one D2H sync per call), so this layout is eager-only.

The per-rank compute is `Glm4MoeMoE(ep_rank, ep_size)`; tests replace it with an oracle-backed
callable to exercise this file's bookkeeping on CPU with the gloo backend.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch import nn


def expert_range(n_experts: int, rank: int, world: int) -> tuple[int, int]:
    per = n_experts // world
    if per * world != n_experts:
        raise ValueError(f"{n_experts} experts do not divide over {world} ranks")
    return rank * per, (rank + 1) * per


def token_slice(T: int, rank: int, world: int) -> tuple[int, int]:
    """Tokens owned by `rank`: contiguous ceil(T/R) chunks (same rule as the C library's
    `with_shared == 2` shared-expert slice)."""
    per = -(-T // world)
    lo = min(rank * per, T)
    return lo, min(lo + per, T)


def dispatch_plan(idx: torch.Tensor, n_experts: int, world: int):
    """For owned tokens with routing `idx [n, k]`: which ranks host at least one selected expert.
    Returns (send_mask [n, R] bool, send_counts [R])."""
    per = n_experts // world
    dest = torch.div(idx, per, rounding_mode="floor")          # [n, k] rank hosting each expert
    mask = torch.zeros((idx.shape[0], world), dtype=torch.bool, device=idx.device)
    mask.scatter_(1, dest, True)
    return mask, mask.sum(0)


class PeerComm:
    """One peer-mapped communication buffer per process for the fused all-reduce + residual add + RMSNorm
    (`csrc/ep_allreduce.cu`): replaces `dist.all_reduce(y)` of the reference's
    `RowParallelLinear.forward` (`nanovllm/layers/linear.py:153-154`) TOGETHER with the add + RMSNorm that follows
    the attention block (`layers/layernorm.py:29-40`, `decode_layer.py:98`).  Collective constructor: every rank
    allocates `glm4moe_ep_comm_bytes(max_tokens, H, dtype, world)` with `glm4moe_p2p_alloc`, the cudaIpc handles
    are exchanged on the host once (`all_gather_object`), every peer's buffer is mapped.  Shared by all layers;
    every rank must make the same sequence of calls with the same token counts."""

    def __init__(self, max_tokens: int, hidden_size: int, dtype: torch.dtype, group=None, device=None):
        import ctypes as C
        from . import _lib
        lib = _lib.load()
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.max_tokens, self.hidden_size, self.dtype = int(max_tokens), int(hidden_size), dtype
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        nbytes = lib.glm4moe_ep_comm_bytes(self.max_tokens, self.hidden_size, _lib.DTYPE_CODE[dtype], self.world)
        if nbytes == 0:
            raise _lib.Glm4MoeLibraryError(f"no peer-memory all-reduce for H={hidden_size}, {dtype}, world={self.world}")
        self.nbytes = nbytes
        ptr, hbuf = C.c_void_p(), C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            _lib.check(lib.glm4moe_p2p_alloc(nbytes + 1024, C.byref(ptr), hbuf), "p2p_alloc")
        self._ptr = ptr.value
        base = (self._ptr + 1023) // 1024 * 1024
        gathered = [None] * self.world
        dist.all_gather_object(gathered, (hbuf.raw, base - self._ptr), group=group)
        self.bases = (C.c_void_p * self.world)()
        self._opened = []
        for r, (h, off) in enumerate(gathered):
            if r == self.rank:
                self.bases[r] = base
                continue
            p = C.c_void_p()
            with torch.cuda.device(self.device):
                _lib.check(lib.glm4moe_p2p_open(h, C.byref(p)), "p2p_open")
            self._opened.append(p.value)
            self.bases[r] = p.value + off
        dist.barrier(group=group)      # nobody launches before every mapping exists

    def allreduce_add_rmsnorm(self, partial: torch.Tensor, residual: torch.Tensor | None = None,
                              weight: torch.Tensor | None = None, eps: float = 1e-6):
        """`y = all_reduce(partial)` (fp32 sum in rank order, one rounding), then the reference's
        `add_rms_forward(y, residual)` / `rms_forward(y)` in the same kernel.  Returns `(out, residual_out)`
        with a residual, `out` without; `weight=None` is a plain all-reduce."""
        from . import _lib
        H = partial.shape[-1]
        x = partial.reshape(-1, H)
        if not x.is_contiguous():
            x = x.contiguous()
        _lib.require_cuda(x, "partial")
        if x.dtype != self.dtype or H != self.hidden_size:
            raise _lib.Glm4MoeLibraryError(f"partial {x.dtype} [.., {H}] != comm {self.dtype} [.., {self.hidden_size}]")
        T = x.shape[0]
        if T > self.max_tokens:
            raise _lib.Glm4MoeLibraryError(f"{T} tokens exceed the comm buffer's {self.max_tokens}")
        r = None
        if residual is not None:
            if weight is None:
                raise _lib.Glm4MoeLibraryError("a residual needs the norm weight")
            r = residual.reshape(-1, H)
            if not r.is_contiguous():
                r = r.contiguous()
            if r.shape != x.shape or r.dtype != x.dtype:
                raise _lib.Glm4MoeLibraryError("residual must match partial in shape and dtype")
        out = torch.empty_like(x)
        res_out = torch.empty_like(x) if r is not None else None
        if T:
            with torch.cuda.device(x.device):
                _lib.check(_lib.load().glm4moe_ep_allreduce_add_rmsnorm(
                    self.bases, self.rank, self.world, self.max_tokens, x.data_ptr(), _lib.ptr(r), _lib.ptr(weight),
                    float(eps), T, H, _lib.DTYPE_CODE[x.dtype], out.data_ptr(), _lib.ptr(res_out),
                    _lib.stream_ptr(x.device)), "ep_allreduce_add_rmsnorm")
        out = out.reshape(partial.shape)
        return out if r is None else (out, res_out.reshape(partial.shape))

    def all_reduce(self, partial: torch.Tensor) -> torch.Tensor:
        return self.allreduce_add_rmsnorm(partial)

    def close(self) -> None:
        from . import _lib
        if self._ptr is None:
            return
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        for p in self._opened:
            _lib.load().glm4moe_p2p_close(p)
        dist.barrier(group=self.group)
        _lib.load().glm4moe_p2p_free(self._ptr)
        self._opened, self._ptr = [], None


class ExpertParallelMoE(nn.Module):
    _p2p_maps: dict = {}             # (group, local allocation) -> mappings of a shared peer workspace

    def __init__(self, local_moe: nn.Module, n_experts: int, group=None, layout: str = "allreduce"):
        super().__init__()
        if layout not in ("allreduce", "alltoall", "p2p"):
            raise ValueError(layout)
        self.local = local_moe
        self.n_experts = n_experts
        self.group = group
        self.layout = layout
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self._peer_bases = None      # ctypes array of every rank's workspace base (layout "p2p")
        self._peer_opened = []
        self._p2p_shared = None      # entry of _p2p_maps when the workspace (and its mappings) are shared

    # ---- layout "p2p": one-time handle exchange ---------------------------------------------------
    def setup_p2p(self, max_tokens: int | None = None) -> None:
        """Collective: allocate / export the local workspace, exchange the cudaIpc handles on the host
        and map every peer's workspace.  Called by the first forward if not done explicitly."""
        import ctypes as C
        from . import _lib
        handle, off = self.local.ipc_handle(max_tokens)
        # blocks that share ONE peer-mapped workspace (Glm4MoeMoE.share_workspace) share its mappings too: the
        # first of them exchanges the handles, the others reuse what it opened.  Every rank builds the same model,
        # so every rank takes this branch for the same blocks and the collective below stays matched.
        shared = type(self)._p2p_maps.get((id(self.group), self.local._ipc_ptr))
        if shared is not None:
            self._peer_bases, self._p2p_shared = shared["bases"], shared
            shared["users"] += 1
            # this block's layer_create has just zeroed the counters of the shared workspace: no rank may start a
            # forward (and push flags into a peer) before every rank is past that point
            dist.barrier(group=self.group)
            return
        gathered = [None] * self.world
        dist.all_gather_object(gathered, (handle, off), group=self.group)
        lib = _lib.load()
        bases = (C.c_void_p * self.world)()
        for r, (h, o) in enumerate(gathered):
            if r == self.rank:
                bases[r] = self.local.workspace_base()
                continue
            ptr = C.c_void_p()
            with torch.cuda.device(self.local.gate.weight.device):
                _lib.check(lib.glm4moe_p2p_open(h, C.byref(ptr)), "p2p_open")
            self._peer_opened.append(ptr.value)
            bases[r] = ptr.value + o
        self._peer_bases = bases
        if getattr(self.local, "_peer_pool_entry", None) is not None:
            self._p2p_shared = {"bases": bases, "opened": self._peer_opened, "users": 1,
                                "key": (id(self.group), self.local._ipc_ptr)}
            type(self)._p2p_maps[self._p2p_shared["key"]] = self._p2p_shared
            self._peer_opened = []
        dist.barrier(group=self.group)     # nobody launches before every mapping exists

    def close_p2p(self) -> None:
        from . import _lib
        shared, self._p2p_shared = self._p2p_shared, None
        if shared is not None:
            shared["users"] -= 1
            if shared["users"] > 0:                 # another block still runs on these mappings
                self._peer_bases = None
                return
            type(self)._p2p_maps.pop(shared["key"], None)
            self._peer_opened = shared["opened"]
        if self._peer_opened:
            torch.cuda.synchronize()
            dist.barrier(group=self.group)
            for p in self._peer_opened:
                _lib.load().glm4moe_p2p_close(p)
        self._peer_opened, self._peer_bases = [], None

    def _forward_p2p(self, h: torch.Tensor) -> torch.Tensor:
        if self._peer_bases is None:
            self.setup_p2p()
        return self.local.forward_ep_p2p(h, self._peer_bases)

    @classmethod
    def build(cls, config, group=None, layout: str = "allreduce", **kw) -> "ExpertParallelMoE":
        from .moe import Glm4MoeMoE
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        local = Glm4MoeMoE(config, ep_rank=rank, ep_size=world, peer_workspace=(layout == "p2p"), **kw)
        return cls(local, config.n_routed_experts, group, layout)

    # ---- layout A --------------------------------------------------------------------------
    def _forward_allreduce(self, h: torch.Tensor) -> torch.Tensor:
        y = self.local(h)                       # local experts + shared expert on this rank's slice
        dist.all_reduce(y, group=self.group)
        return y

    # ---- layout B --------------------------------------------------------------------------
    def _forward_alltoall(self, h: torch.Tensor) -> torch.Tensor:
        T, H = h.shape
        R, r = self.world, self.rank
        lo, hi = token_slice(T, r, R)
        own = h[lo:hi]
        n_own = hi - lo
        # routing of the owned tokens decides the destinations (self is always a destination: the
        # shared expert of an owned token runs on its owner)
        if n_own:
            idx, _ = self.local.gate(own)
            mask, _ = dispatch_plan(idx, self.n_experts, R)
            mask[:, r] = True
        else:
            mask = torch.zeros((0, R), dtype=torch.bool, device=h.device)
        send_counts = mask.sum(0)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts, group=self.group)
        sc, rc = send_counts.tolist(), recv_counts.tolist()          # host sync
        # rows ordered by destination rank, token-ascending inside a destination
        tok_of_row = torch.cat([torch.nonzero(mask[:, d]).flatten() for d in range(R)]) if n_own else \
            torch.zeros(0, dtype=torch.long, device=h.device)
        send = own[tok_of_row]
        recv = torch.empty((sum(rc), H), dtype=h.dtype, device=h.device)
        dist.all_to_all_single(recv, send, rc, sc, group=self.group)
        # put this rank's own tokens first so the shared expert covers exactly rows [0, n_own)
        starts = [sum(rc[:d]) for d in range(R)]
        order = list(range(starts[r], starts[r] + rc[r])) + \
            [i for d in range(R) if d != r for i in range(starts[d], starts[d] + rc[d])]
        order_t = torch.tensor(order, dtype=torch.long, device=h.device)
        part = self.local(recv[order_t], shared_range=(0, rc[r]))    # local experts (+ shared on own rows)
        back = torch.empty_like(part)
        back[order_t] = part
        ret = torch.empty_like(send)
        dist.all_to_all_single(ret, back, sc, rc, group=self.group)
        acc = torch.zeros((n_own, H), dtype=torch.float32, device=h.device)
        acc.index_add_(0, tok_of_row, ret.float())
        mine = acc.to(h.dtype)
        per = -(-T // R)
        padded = torch.zeros((per, H), dtype=h.dtype, device=h.device)
        padded[:n_own] = mine
        gathered = torch.empty((R * per, H), dtype=h.dtype, device=h.device)
        dist.all_gather_into_tensor(gathered, padded, group=self.group)
        return gathered[:T].contiguous()

    def forward_prenorm(self, partial: torch.Tensor, residual: torch.Tensor | None, norm_weight: torch.Tensor,
                        eps: float, comm: "PeerComm | None" = None):
        """`decode_layer.py:96-99` from the o_proj output to the block's output: `partial` is this rank's o_proj
        result BEFORE the all-reduce (`linear.py:153-154`).  Layout "p2p" with a `PeerComm`: one library call
        (fused all-reduce + add + norm kernel, then the expert-parallel block).  Otherwise the same steps with
        NCCL: all-reduce, `glm4moe_add_rmsnorm`, this module's `forward`.  Returns `(out, residual_out)`."""
        comm = comm if comm is not None else getattr(self, "comm", None)
        if self.layout == "p2p" and comm is not None:
            if self._peer_bases is None:
                self.setup_p2p()
            return self.local.forward_ep_p2p_prenorm(partial, residual, norm_weight, eps, self._peer_bases, comm)
        rmsnorm = getattr(self, "norm_fn", None)        # tests on CPU inject the oracle's norm
        if rmsnorm is None:
            from .layernorm import rmsnorm
        y = partial.clone()
        dist.all_reduce(y, group=self.group)
        normed = rmsnorm(y, norm_weight, eps, residual)
        if residual is None:
            return self.forward(normed), None
        return self.forward(normed[0]), normed[1]

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        h = hidden_states.reshape(-1, hidden_states.shape[-1])
        if self.layout == "p2p":
            return self._forward_p2p(h)
        if self.layout == "allreduce":
            return self._forward_allreduce(h)
        return self._forward_alltoall(h)
