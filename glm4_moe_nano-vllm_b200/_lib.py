"""ctypes binding of `csrc/libglm4moe_b200.so` (header: include/glm4_moe_b200.h).

There is no CPU fallback: if the library is missing or fails to load, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .build import LIB_PATH

ABI_VERSION = 2          # GLM4MOE_ABI_VERSION of include/glm4_moe_b200.h
F32, F16, BF16 = 0, 1, 2
IMPL_AUTO, IMPL_SIMT, IMPL_UMMA = 0, 1, 2
W_NATIVE, W_FP8_BLOCK128, W_FP8_BLOCK128_POW2 = 0, 1, 2

DTYPE_CODE = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}

# every symbol include/glm4_moe_b200.h declares (tests check the export list against the header)
SYMBOLS = [
    "glm4moe_abi_version", "glm4moe_debug_set", "glm4moe_error_string", "glm4moe_router_logits", "glm4moe_router_select",
    "glm4moe_route_small", "glm4moe_permute", "glm4moe_permute_static", "glm4moe_permute_scratch_elems", "glm4moe_tile_capacity", "glm4moe_gather_rows",
    "glm4moe_expert_gate_up_simt", "glm4moe_expert_down_simt", "glm4moe_mlp_forward", "glm4moe_combine",
    "glm4moe_workspace_bytes", "glm4moe_layer_create", "glm4moe_layer_destroy",
    "glm4moe_layer_forward", "glm4moe_layer_forward_ex", "glm4moe_layer_forward_timed", "glm4moe_layer_forward_host", "glm4moe_layer_view",
    "glm4moe_layer_last_launches", "glm4moe_layer_last_impl", "glm4moe_layer_last_fused", "glm4moe_layer_tile_rows",
    "glm4moe_p2p_alloc", "glm4moe_p2p_open", "glm4moe_p2p_close", "glm4moe_p2p_free", "glm4moe_layer_forward_ep_p2p", "glm4moe_set_ep_sliced", "glm4moe_add_rmsnorm",
    "glm4moe_layer_forward_prenorm", "glm4moe_ep_comm_bytes", "glm4moe_ep_allreduce_add_rmsnorm",
    "glm4moe_layer_forward_ep_p2p_prenorm", "glm4moe_ep_last_error", "glm4moe_ep_set_timeout_ms",
]


class Config(C.Structure):
    """Mirror of `glm4moe_config_t`."""
    _fields_ = [
        ("hidden_size", C.c_int32), ("n_routed_experts", C.c_int32),
        ("num_experts_per_tok", C.c_int32), ("moe_intermediate_size", C.c_int32),
        ("routed_scaling_factor", C.c_float), ("dtype", C.c_int32), ("weight_format", C.c_int32),
        ("max_tokens", C.c_int32), ("expert_begin", C.c_int32), ("expert_end", C.c_int32),
        ("with_shared", C.c_int32), ("impl", C.c_int32),
    ]


class Glm4MoeLibraryError(RuntimeError):
    pass


_lib = None


def load() -> C.CDLL:
    """Load the CUDA library; raise loudly when it is not built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Glm4MoeLibraryError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). The GLM4-MoE path has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, f32, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    lib.glm4moe_abi_version.restype = i32
    lib.glm4moe_debug_set.argtypes = [i32, i32]
    lib.glm4moe_error_string.restype = C.c_char_p
    lib.glm4moe_error_string.argtypes = [i32]
    lib.glm4moe_router_logits.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp]
    lib.glm4moe_router_select.argtypes = [vp, vp, i32, i32, i32, f32, i32, vp, vp, vp]
    lib.glm4moe_route_small.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, f32, i32, i32, i32, i32, i32, i32,
                                        vp, vp, vp, vp, vp, vp, vp, i32, vp, vp, vp, vp, i32, vp]
    lib.glm4moe_permute.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, vp, vp, vp]
    lib.glm4moe_permute_static.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, vp, vp]
    lib.glm4moe_permute_scratch_elems.restype = sz
    lib.glm4moe_permute_scratch_elems.argtypes = [i32, i32]
    lib.glm4moe_tile_capacity.restype = C.c_int32
    lib.glm4moe_tile_capacity.argtypes = [i32, i32, i32, i32]
    lib.glm4moe_gather_rows.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]
    lib.glm4moe_expert_gate_up_simt.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]
    lib.glm4moe_expert_down_simt.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]
    lib.glm4moe_mlp_forward.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp]
    lib.glm4moe_combine.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]
    lib.glm4moe_workspace_bytes.restype = sz
    lib.glm4moe_workspace_bytes.argtypes = [C.POINTER(Config)]
    lib.glm4moe_layer_create.argtypes = [C.POINTER(Config), vp, vp, vp, vp, vp, vp, vp, sz, C.POINTER(vp)]
    lib.glm4moe_layer_destroy.argtypes = [vp]
    lib.glm4moe_layer_forward.argtypes = [vp, vp, vp, i32, vp]
    lib.glm4moe_layer_forward_ex.argtypes = [vp, vp, vp, i32, i32, i32, vp]
    lib.glm4moe_layer_forward_timed.argtypes = [vp, vp, vp, i32, vp, C.POINTER(C.c_float)]
    lib.glm4moe_layer_forward_host.argtypes = [vp, vp, vp, i32, vp, vp, vp]
    lib.glm4moe_layer_view.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(sz)]
    lib.glm4moe_layer_last_launches.argtypes = [vp]
    lib.glm4moe_layer_last_impl.argtypes = [vp]
    lib.glm4moe_layer_last_fused.argtypes = [vp]
    lib.glm4moe_layer_tile_rows.argtypes = [vp, i32]
    lib.glm4moe_p2p_alloc.argtypes = [sz, C.POINTER(vp), C.c_char_p]
    lib.glm4moe_p2p_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.glm4moe_p2p_close.argtypes = [vp]
    lib.glm4moe_add_rmsnorm.argtypes = [vp, vp, vp, f32, i32, i32, i32, vp, vp, vp]
    lib.glm4moe_p2p_free.argtypes = [vp]
    lib.glm4moe_layer_forward_ep_p2p.argtypes = [vp, vp, vp, i32, C.POINTER(vp), i32, i32, vp]
    lib.glm4moe_set_ep_sliced.argtypes = [i32]
    lib.glm4moe_layer_forward_prenorm.argtypes = [vp, vp, vp, vp, f32, vp, vp, i32, vp]
    lib.glm4moe_ep_comm_bytes.restype = sz
    lib.glm4moe_ep_comm_bytes.argtypes = [i32, i32, i32, i32]
    lib.glm4moe_ep_allreduce_add_rmsnorm.argtypes = [C.POINTER(vp), i32, i32, i32, vp, vp, vp, f32, i32, i32, i32,
                                                     vp, vp, vp]
    lib.glm4moe_layer_forward_ep_p2p_prenorm.argtypes = [vp, vp, vp, vp, f32, vp, vp, i32, C.POINTER(vp),
                                                         C.POINTER(vp), i32, i32, i32, vp]
    lib.glm4moe_ep_last_error.argtypes = [C.POINTER(C.c_int32), i32]
    lib.glm4moe_ep_set_timeout_ms.argtypes = [i32]
    for name in SYMBOLS:
        if lib.glm4moe_abi_version is None or not hasattr(lib, name):
            raise Glm4MoeLibraryError(f"{LIB_PATH} does not export {name}")
    if lib.glm4moe_abi_version() != ABI_VERSION:
        raise Glm4MoeLibraryError("ABI version mismatch between _lib.py and the built library")
    _lib = lib
    return lib


NVTX = os.environ.get("GLM4MOE_NVTX", "0") == "1"


class nvtx_range:
    """`with nvtx_range("glm4moe.forward", T):` — an NVTX range around a host entry point when GLM4MOE_NVTX=1
    (the reference has no tracing of its own; nsys / ncu --nvtx pick these up), else free."""

    __slots__ = ("on",)

    def __init__(self, name: str, T: int | None = None):
        self.on = NVTX
        if self.on:
            torch.cuda.nvtx.range_push(name if T is None else f"{name} T={T}")

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        if self.on:
            torch.cuda.nvtx.range_pop()
        return False


def ep_last_error(clear: bool = False) -> dict | None:
    """What a cross-GPU wait that gave up recorded (glm4moe_ep_last_error), or None.  Makes no CUDA call: usable
    after the trap that follows such a failure has poisoned the context."""
    lib = load()
    info = (C.c_int32 * 6)()
    code = lib.glm4moe_ep_last_error(info, int(clear))
    if code == 0:
        return None
    return {"code": code, "error": lib.glm4moe_error_string(code).decode(), "rank": info[1], "waited_for_call": info[2],
            "last_flag": info[3], "site": {1: "expert combine", 2: "all-reduce + norm"}.get(info[4], str(info[4])),
            "peer": info[5]}


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = load().glm4moe_error_string(code).decode()
        if code > 0:                      # a cudaError_t: say so when an expert-parallel wait is behind it
            ep = ep_last_error()
            if ep is not None:
                msg += f" [{ep['error']}: rank {ep['rank']} waited for call {ep['waited_for_call']} of rank " \
                       f"{ep['peer']} in the {ep['site']}, last flag {ep['last_flag']}]"
        raise Glm4MoeLibraryError(f"{what or 'glm4moe call'} failed: {msg} (code {code})")


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise Glm4MoeLibraryError(f"{name} must live on a CUDA device (got {t.device}); "
                                  "the GLM4-MoE path has no CPU implementation")
    if not t.is_contiguous():
        raise Glm4MoeLibraryError(f"{name} must be contiguous")
