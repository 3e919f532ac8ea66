"""`glm4moe_permute_static`: the permutation for hosts that cannot give the library any scratch memory.

Some users of this library never get to allocate on the GPU: a plugin that its runtime hands nothing but the
input and output buffers of the op, a process whose device memory was fully handed out (KV cache, other tenants)
before this library was loaded, a call site where cudaMalloc is off limits.  For them `glm4moe_permute_static`
keeps the per-block count table of the permutation in a fixed array inside the library's own device image
(limit: T <= GLM4MOE_PERMUTE_STATIC_MAX_TOKENS = 16384) and takes no `scratch` argument.

What is checked here, through the C ABI: the new entry point returns exactly what `glm4moe_permute` returns with
caller scratch -- the permutation is integer work, so the tolerance is the one every permutation test in this
repo uses: bit-exact -- on two calls in a row (different inputs, then the first inputs again), so that whatever
one call leaves in the shared array is shown not to leak into the next; and that a T beyond the limit is refused
with an error instead of overrunning the array.  The file sorts last on purpose (`pytest -x` stops at the first
failure; a failure here must not hide the parity tests), and passes alone in a fresh process as well as after
the rest of the suite: it assumes nothing about the array's contents.
"""
import pytest
import torch

from glm4_moe_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_TOKENS = 16384      # GLM4MOE_PERMUTE_STATIC_MAX_TOKENS in include/glm4_moe_b200.h
TILE_ROWS = 16


def _routing(T, E, k, seed):
    """k distinct experts per token, ascending (what the router emits), plus fp32 routing weights."""
    g = torch.Generator().manual_seed(seed)
    idx = torch.rand((T, E), generator=g).topk(k, dim=1).indices.sort(dim=1).values.to(torch.int32)
    w = torch.rand((T, k), generator=g)
    return idx.cuda(), w.cuda()


def _outputs(T, E, k, cap, fill):
    dev = "cuda"
    return {
        "off": torch.full((E + 1,), fill, dtype=torch.int32, device=dev),
        "row_token": torch.full((T * k,), fill, dtype=torch.int32, device=dev),
        "row_weight": torch.full((T * k,), float(fill), dtype=torch.float32, device=dev),
        "slot_row": torch.full((T * k,), fill, dtype=torch.int32, device=dev),
        "tiles": torch.full((cap, 4), fill, dtype=torch.int32, device=dev),
        "num_tiles": torch.full((4,), fill, dtype=torch.int32, device=dev),
    }


def _permute(lib, idx, w, T, E, k, e_lo, e_hi, sh_lo, sh_hi, static):
    cap = lib.glm4moe_tile_capacity(T, k, E, TILE_ROWS)
    o = _outputs(T, E, k, cap, -7 if static else -9)
    head = (idx.data_ptr(), w.data_ptr(), T, k, E, e_lo, e_hi, TILE_ROWS, sh_lo, sh_hi, o["off"].data_ptr(),
            o["row_token"].data_ptr(), o["row_weight"].data_ptr(), o["slot_row"].data_ptr(), o["tiles"].data_ptr(),
            cap, o["num_tiles"].data_ptr())
    if static:
        _lib.check(lib.glm4moe_permute_static(*head, _lib.stream_ptr()))
    else:
        # the caller's scratch is poisoned: neither entry point may depend on what the table held before
        scratch = torch.full((lib.glm4moe_permute_scratch_elems(T, E),), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
        _lib.check(lib.glm4moe_permute(*head, scratch.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    return {n: t.cpu() for n, t in o.items()}


def _same(a, b, T, k):
    n_rows = int(b["off"][-1])
    n_tiles = int(b["num_tiles"][0])
    assert n_tiles > 0
    assert torch.equal(a["off"], b["off"])
    assert int(a["num_tiles"][0]) == n_tiles
    assert torch.equal(a["slot_row"], b["slot_row"])
    # rows past the routed count and tiles past the list are not written by either entry point
    assert torch.equal(a["row_token"][:n_rows], b["row_token"][:n_rows])
    assert torch.equal(a["row_weight"][:n_rows], b["row_weight"][:n_rows])
    assert torch.equal(a["tiles"][:n_tiles], b["tiles"][:n_tiles])
    assert n_rows <= T * k


# (T, E, k, e_lo, e_hi, shared_lo, shared_hi): the three-kernel path (T > 256, the one that uses the array) at a
# ragged size, the expert-parallel filter, the largest T the array holds, and the single-block path for completeness
CASES = [
    (300, 160, 8, 0, 160, 0, 300),
    (4096, 128, 8, 64, 128, 0, 0),
    (MAX_TOKENS, 256, 8, 0, 256, 0, MAX_TOKENS),
    (33, 32, 8, 0, 32, 0, 33),
]


@pytest.mark.parametrize("T,E,k,e_lo,e_hi,sh_lo,sh_hi", CASES)
def test_permute_static_matches_permute_on_two_calls_in_a_row(T, E, k, e_lo, e_hi, sh_lo, sh_hi):
    lib = _lib.load()
    first = _routing(T, E, k, seed=T * 131 + E)
    second = _routing(T, E, k, seed=T * 131 + E + 1)
    want_first = _permute(lib, *first, T, E, k, e_lo, e_hi, sh_lo, sh_hi, static=False)
    want_second = _permute(lib, *second, T, E, k, e_lo, e_hi, sh_lo, sh_hi, static=False)
    assert not torch.equal(want_first["slot_row"], want_second["slot_row"])   # the two calls do differ
    # call 1, call 2 on other routing (the array now holds call 1's counts), then call 1's inputs once more
    _same(_permute(lib, *first, T, E, k, e_lo, e_hi, sh_lo, sh_hi, static=True), want_first, T, k)
    _same(_permute(lib, *second, T, E, k, e_lo, e_hi, sh_lo, sh_hi, static=True), want_second, T, k)
    _same(_permute(lib, *first, T, E, k, e_lo, e_hi, sh_lo, sh_hi, static=True), want_first, T, k)


def test_permute_static_after_a_different_shape():
    """The array is shared by every call: a large call followed by a smaller one with another expert count (another
    row stride in the table) must not see the large call's leftovers."""
    lib = _lib.load()
    big = (8192, 256, 8)
    small = (1000, 96, 4)
    for T, E, k in (big, small, big):
        a = _routing(T, E, k, seed=T + E)
        want = _permute(lib, *a, T, E, k, 0, E, 0, T, static=False)
        _same(_permute(lib, *a, T, E, k, 0, E, 0, T, static=True), want, T, k)


def test_permute_static_refuses_more_tokens_than_the_array_holds():
    lib = _lib.load()
    T, E, k = MAX_TOKENS + 1, 128, 8
    idx, w = _routing(T, E, k, seed=5)
    cap = lib.glm4moe_tile_capacity(T, k, E, TILE_ROWS)
    o = _outputs(T, E, k, cap, -7)
    rc = lib.glm4moe_permute_static(idx.data_ptr(), w.data_ptr(), T, k, E, 0, E, TILE_ROWS, 0, T, o["off"].data_ptr(),
                                    o["row_token"].data_ptr(), o["row_weight"].data_ptr(), o["slot_row"].data_ptr(),
                                    o["tiles"].data_ptr(), cap, o["num_tiles"].data_ptr(), _lib.stream_ptr())
    torch.cuda.synchronize()
    assert rc == -3                                         # GLM4MOE_ERR_WORKSPACE
    assert b"workspace" in lib.glm4moe_error_string(rc).lower()
    with pytest.raises(_lib.Glm4MoeLibraryError):
        _lib.check(rc)
    assert bool((o["slot_row"] == -7).all()) and bool((o["off"] == -7).all())   # nothing was launched
