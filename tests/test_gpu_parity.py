"""GPU parity tests proper: every call goes through the C ABI (ctypes) into the sm_100a kernels and
is compared with the CPU oracle (oracle/moe_oracle.py) and the golden outputs of the real reference
(tests/golden).  Bars (BASELINE.json north_star): expert indices and the permutation bit-exact;
layer outputs within rtol 2e-2 of the fp32 truth (|d| <= rtol*|ref| + rtol*rms(ref))
and within the reference's own rtol=atol=1e-2 of its native output (nanovllm/tests/moe.py:40)."""
import os

import pytest
import torch

import glm4_moe_b200
from glm4_moe_b200 import synth, _lib
from glm4_moe_b200 import layernorm as LN
from oracle import moe_oracle as O
from conftest import load_golden, dtype_of

pytestmark = pytest.mark.gpu

RTOL_BF16 = 2e-2     # north_star: bf16 outputs vs the fp32 reference
FULL = ["tiny_fp32_T64", "mini_bf16_T48", "mini_fp16_T33", "mini_bf16_T1", "small_bf16_T256"]


def _cfg(shape):
    from transformers.models.glm4_moe import Glm4MoeConfig
    return Glm4MoeConfig(hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                         moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                         routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)


def _module(shape, dt, weights, **kw):
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dt)
    try:
        moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape), **kw)
    finally:
        torch.set_default_dtype(prev)
    moe.load_stacked(weights)
    return moe.cuda()


def _close(got, truth, rtol):
    got, truth = got.float().cpu(), truth.float().cpu()
    rms = truth.pow(2).mean().sqrt()
    bad = (got - truth).abs() > rtol * truth.abs() + rtol * rms
    rel = ((got - truth).norm() / truth.norm()).item()
    return int(bad.sum()), rel


def _ulp(dt):
    return {torch.bfloat16: 2.0 ** -8, torch.float16: 2.0 ** -11, torch.float32: 2.0 ** -24}[dt]


def _impls(shape, dt):
    ok = dt in (torch.bfloat16, torch.float16) and shape.H % 128 == 0 and shape.I % 128 == 0
    return ["simt", "umma"] if ok else ["simt"]


def _log_routing(kind, shape, dt, tokens, flips, against):
    """The observed count, not only the cap: one JSON line per routing comparison into gpurun_out/routing_flips.jsonl
    (summarised in profiles/).  `flips` = tokens whose expert SET differs from `against`, each one already explained
    as a k-th / (k+1)-th boundary pair by the caller."""
    import json
    rec = {"test": os.environ.get("PYTEST_CURRENT_TEST", "").split(" ")[0], "kind": kind, "shape": shape.name,
           "dtype": str(dt), "tokens": int(tokens), "boundary_flips": int(flips), "against": against}
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    try:
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "routing_flips.jsonl"), "a") as f:
            f.write(json.dumps(rec) + "\n")
    except OSError:
        pass
    print("routing:", json.dumps(rec))


def _explain_routing(idx, logits, x, gate_w, bias, shape, dt, max_frac=2e-3):
    """Kernel routing (`idx` canonical [T,k], `logits` the kernel's own) against the oracle router run on
    this machine's CPU.  Index sets must be equal except on tokens where (a) the two sets differ in exactly
    one expert pair, (b) that pair sits on the k-th / (k+1)-th boundary within two model-dtype ulps of
    `scores_for_choice`, and (c) the two logit computations actually rounded one of the pair's logits
    differently (fp32 accumulation order).  Returns the number of such tokens."""
    ref_logits = O.router_logits(x, gate_w)
    # ties in scores_for_choice: torch.topk's pick is implementation-defined, the kernel's rule is pinned
    # (lower expert id), so the oracle is asked for that rule
    oi, ow = O.canonical_routing(*O.select_from_logits(ref_logits, bias, shape.k, shape.routed_scaling_factor,
                                                       tie_rule="lowest_index"))
    mism = (idx != oi).any(-1)
    d = (logits.float() - ref_logits.float()).abs()
    sc = logits.float().sigmoid().to(dt).float() + bias
    for t in torch.nonzero(mism).flatten().tolist():
        only = list(set(idx[t].tolist()) ^ set(oi[t].tolist()))
        assert len(only) == 2, f"token {t}: sets differ in more than one expert"
        gap = abs(float(sc[t, only[0]] - sc[t, only[1]]))
        assert gap <= 2 * _ulp(dt), f"token {t}: real routing mismatch (gap {gap})"
        if float(d[t, only].max()) == 0:
            # identical logits: then the model-dtype rounding of sigmoid (moe.py:37) must be what differs, i.e.
            # one of the two exact sigmoids sits on a rounding boundary of the model dtype to within the
            # accuracy of an fp32 expf (torch's vectorised CPU sigmoid vs the kernel's 1 / (1 + expf(-x)))
            assert dt != torch.float32, f"token {t}: mismatch without a logit flip"
            u = _ulp(dt)
            on_boundary = False
            for e in only:
                v = float(torch.sigmoid(logits[t, e].double()))
                ulp_v = u * 2.0 ** (torch.tensor(v).log2().floor().item() + 1)      # spacing of dt at v
                frac = (v / ulp_v) % 1.0
                on_boundary |= abs(frac - 0.5) < 2e-3
            assert on_boundary, (f"token {t}: mismatch without a logit flip or a sigmoid rounding boundary: experts "
                                 f"{only}, logits {logits[t, only].tolist()}, scores_for_choice "
                                 f"{sc[t, only].tolist()}, bias {bias[only].tolist()}")
    _log_routing("layer", shape, dt, idx.shape[0], mism.sum(), "oracle router on this host's CPU")
    # observed on B200: at most 1 token in 4096 (profiles/r3_routing_flips.md); the cap leaves room for one token in
    # a small batch and 8x the observed rate in a large one
    assert int(mism.sum()) <= max(1, int(max_frac * idx.shape[0])), (int(mism.sum()), idx.shape[0])
    return int(mism.sum())


def _sample_tokens(idx, E, tile_rows, n_extra=48, seed=0):
    """Tokens whose rows sit where a tile kernel can go wrong: for every expert the first row, the last row
    (ragged tail tile) and both sides of the first tile boundary; the first / last token of the batch;
    plus a few random ones."""
    T = idx.shape[0]
    rt, _, offs = O.permutation(idx, E)
    rt, offs = rt.tolist(), offs.tolist()
    toks = {0, T - 1}
    for e in range(E):
        a, b = offs[e], offs[e + 1]
        if b > a:
            toks.update((rt[a], rt[b - 1]))
            if b - a > tile_rows:
                toks.update((rt[a + tile_rows - 1], rt[a + tile_rows]))
    # (rows ascend by token inside an expert, so the structural picks cluster at both ends of the batch)
    # ... and random tokens from the whole batch until at least `n_extra` of those and 320 tokens in total
    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(T, generator=g).tolist()
    target = min(T, max(320, len(toks) + n_extra))
    for t in perm:
        if len(toks) >= target:
            break
        toks.add(t)
    return torch.tensor(sorted(toks), dtype=torch.long)


def _check_sampled_rows(out, x, wts, shape, dt, idx, w, tile_rows, seed=0, n_extra=48):
    """The MoE is per token: the fp32 truth (and the reference's native-dtype result) of a few hundred sampled
    tokens under the kernel's (already verified) routing, computed by the oracle on the host CPU, against
    the same rows of a full-size batch.  rtol 2e-2 (+ rtol * rms) vs fp32, rtol = atol = 1e-2
    (nanovllm/tests/moe.py:40) vs the native-dtype evaluation."""
    sel = _sample_tokens(idx, shape.E, tile_rows, n_extra, seed)
    xs, ri, rw = x[sel], idx[sel], w[sel]
    with torch.inference_mode():
        truth = O.moe_forward_fp32(xs, wts, shape.k, shape.routed_scaling_factor, routing=(ri, rw))["out"]
        native = O.moe_forward(xs, wts, shape.k, shape.routed_scaling_factor, routing=(ri, rw))["out"]
    got = out[sel.to(out.device)]
    nbad, rel = _close(got, truth, RTOL_BF16)
    assert nbad == 0 and rel < 6e-3, (nbad, rel, int(sel.numel()))
    # The reference's own tolerance (rtol = atol = 1e-2, tests/moe.py:40) against its native-dtype result.  At
    # these shapes outputs reach |y| ~ 2..8 (routed_scaling_factor 2.5: ~7), where one bf16 ulp is 0.016..0.031,
    # so that tolerance is 1-2 ulps and the reference's k model-dtype accumulations (moe.py:151) alone put a
    # fraction o_nt of ITS elements outside the same tolerance of its own fp32 evaluation (measured on the
    # CPU: Air 1.8e-4, GLM-4.5 1.4e-2; an ideal kernel = the fp32 truth rounded once = lands 1.6e-2 of the
    # GLM-4.5 elements outside the native result).  Bar, relative to that measured noise of the reference
    # itself: vs the native result at most 2 * o_nt + 1e-3 of the elements outside (measured Air 5.7e-4,
    # GLM-4.5 2.4e-2), and vs the fp32 truth no more elements outside than the native result has (+1e-3);
    # none further than 4 model-dtype ulps of the value range.
    g, n_, t_ = got.float().cpu(), native.float(), truth.float()

    def outside(a, b):
        return ((a - b).abs() > 1e-2 + 1e-2 * b.abs()).float().mean().item()
    o_nt, o_gn, o_gt = outside(n_, t_), outside(g, n_), outside(g, t_)
    assert o_gn <= 2 * o_nt + 1e-3, (o_gn, o_nt)
    assert o_gt <= o_nt + 1e-3, (o_gt, o_nt)
    assert float((g - n_).abs().max()) <= 4 * _ulp(dt) * float(n_.abs().max()), float((g - n_).abs().max())
    return int(sel.numel()), rel


class _knobs:
    """glm4moe_debug_set for the duration of a block; restores the defaults."""
    DEFAULTS = {4: 0, 5: 1, 6: 0, 10: 1}   # BK32 off, DOWN2 on, tile rows by mean rows per expert, tail boxes on

    def __init__(self, **kw):
        names = {"bk32": 4, "down2": 5, "tile_rows": 6, "tail_box": 10}
        self.values = {names[k_]: v for k_, v in kw.items()}

    def __enter__(self):
        for k_, v in self.values.items():
            _lib.check(_lib.load().glm4moe_debug_set(k_, v))

    def __exit__(self, *a):
        for k_, v in self.DEFAULTS.items():
            _lib.check(_lib.load().glm4moe_debug_set(k_, v))


# ------------------------------------------------------------------------------------------
# router
# ------------------------------------------------------------------------------------------
def _check_router(x, gate_w, bias, shape, dt, ref_idx, ref_w):
    """Three-stage router parity: logits within one model-dtype ulp of the
    oracle's; selection bit-exact given the kernel's own logits; and end-to-end index sets equal
    to the reference's except where a logit rounding flip sits on the k-th/k+1-th boundary."""
    cfg = _cfg(shape)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dt)
    try:
        gate = glm4_moe_b200.GlmMoeSelectTopk(cfg)
    finally:
        torch.set_default_dtype(prev)
    gate.weight.data.copy_(gate_w)
    gate.e_score_correction_bias.copy_(bias)
    gate = gate.cuda()
    idx, w, logits = gate(x.cuda(), return_logits=True)
    idx, w, logits = idx.cpu(), w.cpu(), logits.cpu()
    assert idx.dtype == torch.int64 and w.dtype == dt
    # (1) logits: fp32 accumulate, one rounding; differences only from accumulation order
    ref_logits = O.router_logits(x, gate_w)
    d = (logits.float() - ref_logits.float()).abs()
    tol = 2 * _ulp(dt) * ref_logits.float().abs().clamp_min(1.0) if dt != torch.float32 else 1e-5
    assert bool((d <= tol).all()), f"router logits differ by more than one ulp: {d.max()}"
    flips = float((d > 0).float().mean())
    assert flips < 5e-3 or dt == torch.float32
    # (2) selection given identical logits: bit-exact indices (as sets, canonical order) + weights
    oi, ow = O.select_from_logits(logits, bias, shape.k, shape.routed_scaling_factor)
    oi, ow = O.canonical_routing(oi, ow)
    assert torch.equal(idx, oi), "top-k indices differ from the oracle on identical logits"
    if dt == torch.float32:
        assert torch.allclose(w, ow, rtol=1e-6, atol=0)
    else:
        assert torch.equal(w, ow), "routing weights differ from the oracle on identical logits"
    # (3) end to end vs the reference's own outputs
    ri, rw = O.canonical_routing(ref_idx.long(), ref_w)
    mism = (idx != ri).any(-1)
    if bool(mism.any()):
        # every disputed token must be explained by a logit flip on a boundary expert
        sc = logits.float().sigmoid().to(dt).float() + bias
        for t in torch.nonzero(mism).flatten().tolist():
            a, b = set(idx[t].tolist()), set(ri[t].tolist())
            only = list(a ^ b)
            assert len(only) == 2, f"token {t}: sets differ in more than one expert"
            gap = abs(float(sc[t, only[0]] - sc[t, only[1]]))
            assert gap <= 2 * _ulp(dt), f"token {t}: real routing mismatch (gap {gap})"
            assert float(d[t, only].max()) > 0, f"token {t}: mismatch without a logit flip"
    _log_routing("router", shape, dt, idx.shape[0], mism.sum(), "golden outputs of the real reference")
    assert float(mism.float().mean()) <= 2e-3
    return int(mism.sum())


@pytest.mark.parametrize("name", ["air_router_bf16_T256", "glm45_router_bf16_T64"])
def test_router_big_shapes_vs_reference(name):
    g = load_golden(name)
    c = g["case"]
    shape, dt = synth.SHAPES[c["shape"]], dtype_of(c["dtype"])
    gate_w, bias = synth.make_router_weights(shape, dt)
    x = synth.make_tokens(shape, c["T"], dt, c["batch_index"])
    _check_router(x, gate_w, bias, shape, dt, g["idx"], g["w"])


@pytest.mark.parametrize("name", FULL)
def test_router_vs_reference(name):
    g = load_golden(name)
    c = g["case"]
    shape, dt = synth.SHAPES[c["shape"]], dtype_of(c["dtype"])
    wts = synth.make_weights(shape, dt)
    x = synth.make_tokens(shape, c["T"], dt, c["batch_index"])
    _check_router(x, wts["gate_weight"], wts["bias"], shape, dt, g["idx"], g["w"])


def test_router_tie_rule_is_pinned():
    """Tie stress: zero bias + constant tokens => all scores equal; duplicated router
    rows => pairwise ties.  The pinned rule is 'higher score, then lower expert id'."""
    shape, dt = synth.AIR, torch.bfloat16
    gate_w, _ = synth.make_router_weights(shape, dt)
    gate_w[1::2] = gate_w[0::2]                      # expert 2j+1 duplicates expert 2j
    bias = torch.zeros(shape.E)
    x = synth.make_tokens(shape, 64, dt, 5)
    x[:8] = 0                                        # all logits 0 -> all scores 0.5
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dt)
    gate = glm4_moe_b200.GlmMoeSelectTopk(_cfg(shape))
    torch.set_default_dtype(prev)
    gate.weight.data.copy_(gate_w)
    gate = gate.cuda()
    idx, w, logits = gate(x.cuda(), return_logits=True)
    idx, w, logits = idx.cpu(), w.cpu(), logits.cpu()
    assert torch.equal(idx[:8], torch.arange(shape.k).expand(8, -1))       # lowest ids win
    assert torch.equal(w[:8].float(), torch.full((8, shape.k), 0.125))
    oi, ow = O.select_from_logits(logits, bias, shape.k, shape.routed_scaling_factor, tie_rule="lowest_index")
    oi, ow = O.canonical_routing(oi, ow)
    assert torch.equal(idx, oi) and torch.equal(w, ow)
    # and it is a valid top-k for torch.topk too: same multiset of selected scores
    ti, _ = O.select_from_logits(logits, bias, shape.k, shape.routed_scaling_factor, tie_rule="torch")
    sc = logits.sigmoid().float()
    assert torch.equal(sc.gather(1, idx).sort(-1)[0], sc.gather(1, ti).sort(-1)[0])


# ------------------------------------------------------------------------------------------
# permutation
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,E,k", [(1, 16, 4), (33, 32, 8), (256, 128, 8), (300, 160, 8), (4096, 128, 8)])
def test_permutation_bit_exact(T, E, k):
    lib = _lib.load()
    g = torch.Generator().manual_seed(T * 131 + E)
    idx = torch.stack([torch.randperm(E, generator=g)[:k].sort()[0] for _ in range(T)]).to(torch.int32)
    w = torch.rand((T, k), generator=g)
    row_token, row_slot, offs = O.permutation(idx.long(), E)
    dev = "cuda"
    tr = 16
    cap = lib.glm4moe_tile_capacity(T, k, E, tr)
    d_idx, d_w = idx.to(dev), w.to(dev)
    d_off = torch.empty(E + 1, dtype=torch.int32, device=dev)
    d_rt = torch.full((T * k,), -7, dtype=torch.int32, device=dev)
    d_rw = torch.empty(T * k, dtype=torch.float32, device=dev)
    d_sr = torch.empty(T * k, dtype=torch.int32, device=dev)
    d_tiles = torch.zeros((cap, 4), dtype=torch.int32, device=dev)
    d_nt = torch.zeros(4, dtype=torch.int32, device=dev)
    d_scr = torch.empty(lib.glm4moe_permute_scratch_elems(T, E), dtype=torch.int32, device=dev)
    _lib.check(lib.glm4moe_permute(d_idx.data_ptr(), d_w.data_ptr(), T, k, E, 0, E, tr, 0, T, d_off.data_ptr(),
                                   d_rt.data_ptr(), d_rw.data_ptr(), d_sr.data_ptr(), d_tiles.data_ptr(), cap,
                                   d_nt.data_ptr(), d_scr.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(d_off.cpu(), offs)
    assert torch.equal(d_rt.cpu(), row_token)
    assert torch.equal(d_rw.cpu(), w[row_token.long(), row_slot.long()])
    sr = d_sr.cpu().view(T, k)
    inv = torch.empty(T * k, dtype=torch.int64)
    inv[row_token.long() * k + row_slot.long()] = torch.arange(T * k)
    assert torch.equal(sr.long().view(-1), inv)
    # tile list covers every routed row exactly once, then the shared rows
    nt = int(d_nt[0])
    tiles = d_tiles.cpu()[:nt]
    covered = torch.zeros(T * k + T, dtype=torch.int32)
    for e, r0, n, _ in tiles.tolist():
        assert 0 < n <= tr
        covered[r0:r0 + n] += 1
        if e < E:
            assert offs[e] <= r0 and r0 + n <= offs[e + 1]
        else:
            assert r0 >= T * k
    assert bool((covered == 1).all())
    # expert-parallel filter: only experts [E/2, E) keep rows, order still stable
    _lib.check(lib.glm4moe_permute(d_idx.data_ptr(), d_w.data_ptr(), T, k, E, E // 2, E, tr, 0, 0, d_off.data_ptr(),
                                   d_rt.data_ptr(), d_rw.data_ptr(), d_sr.data_ptr(), d_tiles.data_ptr(), cap,
                                   d_nt.data_ptr(), d_scr.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    keep = idx.long().view(-1)[row_token.long() * k + row_slot.long()] >= E // 2
    n_keep = int(keep.sum())
    assert int(d_off[E]) == n_keep
    assert torch.equal(d_rt.cpu()[:n_keep], row_token[keep])
    assert bool((d_sr.cpu().view(T, k)[idx < E // 2] == -1).all())


# ------------------------------------------------------------------------------------------
# whole block vs the reference's outputs (golden) and the fp32 truth
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", FULL)
def test_layer_vs_reference_golden(name):
    g = load_golden(name)
    c = g["case"]
    shape, dt = synth.SHAPES[c["shape"]], dtype_of(c["dtype"])
    wts = synth.make_weights(shape, dt)
    x = synth.make_tokens(shape, c["T"], dt, c["batch_index"])
    ri, rw = O.canonical_routing(g["idx"].long(), g["w"])
    with torch.inference_mode():
        truth = O.moe_forward_fp32(x, wts, shape.k, shape.routed_scaling_factor, routing=(ri, rw))["out"]
    for impl in _impls(shape, dt):
        moe = _module(shape, dt, wts, impl=impl)
        with torch.inference_mode():
            out = moe(x.cuda())
        torch.cuda.synchronize()
        assert out.shape == (c["T"], shape.H) and out.dtype == dt
        assert moe.last_impl == impl
        idx = moe.workspace_view("idx", c["T"]).cpu().long()
        assert torch.equal(idx, ri), f"{impl}: routing differs from the reference"
        rt, _, offs = O.permutation(ri, shape.E)
        assert torch.equal(moe.workspace_view("expert_offsets").cpu(), offs)
        assert torch.equal(moe.workspace_view("row_token").cpu()[:c["T"] * shape.k], rt)
        rtol = 1e-5 if dt == torch.float32 else RTOL_BF16
        nbad, rel = _close(out, truth, rtol)
        assert nbad == 0, f"{impl}: {nbad} elements outside rtol {rtol} of the fp32 truth (rel-F {rel})"
        assert rel < (1e-5 if dt == torch.float32 else 6e-3)
        # the reference's own tolerance against its native-dtype output (tests/moe.py:40)
        torch.testing.assert_close(out.cpu().float(), g["out"].float(), rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("shape_name,T", [("mini", 48), ("small", 256)])
def test_layer_vs_reference_torch_path_on_the_same_gpu(shape_name, T):
    """SURVEY 8c "O-native": the reference's torch ops (moe.py:28-50,128-176 as restated in the oracle)
    executed on the SAME device in the model dtype — cuBLAS `F.linear`, CUDA `torch.topk`, `index_add_` —
    against the CUDA path.  Index sets must agree except where the two logit computations round a
    boundary logit differently; outputs of tokens with identical routing must agree within the
    reference's own tolerance (tests/moe.py:40) and within rtol 2e-2 (+ rtol * rms)."""
    shape, dt = synth.SHAPES[shape_name], torch.bfloat16
    wts = synth.make_weights(shape, dt)
    x = synth.make_tokens(shape, T, dt, 3)
    dev_w = {k_: v.cuda() for k_, v in wts.items()}
    moe = _module(shape, dt, wts)
    with torch.inference_mode():
        out = moe(x.cuda())
        ref = O.moe_forward(x.cuda(), dev_w, shape.k, shape.routed_scaling_factor)
        ref_logits = O.router_logits(x.cuda(), dev_w["gate_weight"])
    torch.cuda.synchronize()
    idx = moe.workspace_view("idx", T).cpu().long()
    logits = moe.workspace_view("logits", T).cpu().float()
    ri, _ = O.canonical_routing(ref["idx"].cpu().long(), ref["w"].cpu())
    mism = (idx != ri).any(-1)
    dl = (logits - ref_logits.cpu().float()).abs()
    for t in torch.nonzero(mism).flatten().tolist():
        only = list(set(idx[t].tolist()) ^ set(ri[t].tolist()))
        assert len(only) == 2, f"token {t}: sets differ in more than one expert"
        assert float(dl[t, only].max()) > 0, f"token {t}: routing differs without a logit rounding difference"
    assert float(mism.float().mean()) <= 0.02
    same = ~mism
    got, want = out.cpu().float()[same], ref["out"].cpu().float()[same]
    torch.testing.assert_close(got, want, rtol=1e-2, atol=1e-2)
    nbad, rel = _close(got, want, RTOL_BF16)
    assert nbad == 0 and rel < 1e-2, (nbad, rel)


def test_simt_and_umma_agree():
    shape, dt = synth.MINI, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    x = synth.make_tokens(shape, 200, dt, 9).cuda()
    outs = {}
    for impl in ("simt", "umma"):
        moe = _module(shape, dt, wts, impl=impl)
        outs[impl] = moe(x).float().cpu()
    nbad, rel = _close(outs["umma"], outs["simt"], 1e-2)
    assert nbad == 0 and rel < 4e-3


def test_edge_cases_and_properties():
    shape, dt = synth.MINI, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    for impl in ("simt", "umma"):
        moe = _module(shape, dt, wts, impl=impl)
        # empty input
        assert moe(torch.empty((0, shape.H), dtype=dt, device="cuda")).shape == (0, shape.H)
        # 3-D input comes back 2-D (moe.py:130,176); input is not modified
        x = synth.make_tokens(shape, 48, dt, 3).cuda()
        x0 = x.clone()
        y = moe(x.view(4, 12, shape.H))
        assert y.shape == (48, shape.H) and torch.equal(x, x0)
        # permutation equivariance and batch-split consistency, bit-exact
        perm = torch.randperm(48, generator=torch.Generator().manual_seed(1)).cuda()
        assert torch.equal(moe(x[perm]), y[perm])
        assert torch.equal(moe(x[:17]), y[:17])
        # run-to-run determinism (no atomics in the combine)
        assert torch.equal(moe(x), y)
        # every token identical -> 8 experts take all rows (several row tiles per expert)
        xs = x[:1].expand(300, -1).contiguous()
        ys = moe(xs)
        assert torch.equal(ys, ys[:1].expand(300, -1))
        offs = moe.workspace_view("expert_offsets").cpu()
        assert int(offs[-1]) == 300 * shape.k and int((offs[1:] - offs[:-1]).max()) == 300
        assert torch.equal(ys[0], moe(x[:1])[0])


def test_cuda_graph_capture_and_host_entry():
    shape, dt = synth.MINI, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    moe = _module(shape, dt, wts)
    x = synth.make_tokens(shape, 16, dt, 4)
    xs = x.cuda()
    eager = moe(xs)                                   # also sizes the workspace before capture
    torch.cuda.synchronize()
    static_in = xs.clone()
    graph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        moe(static_in)
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(graph):
        static_out = moe(static_in)
    static_in.copy_(synth.make_tokens(shape, 16, dt, 5).cuda())
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(static_out, moe(static_in))
    static_in.copy_(xs)
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(static_out, eager)
    # host-buffer entry point == device entry point
    out_host = moe.forward_host(x.pin_memory())
    assert torch.equal(out_host, eager.cpu())
    assert moe.last_launches == 2      # fused route (logits+select+permute), fused expert MLP + combine



# ------------------------------------------------------------------------------------------
# decode front-end: one fused launch == the three stand-alone kernels; direct mode == general path
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,shape_name,dt", [(1, "glm-4.5-air", torch.bfloat16), (7, "glm-4.5-air", torch.bfloat16),
                                             (32, "glm-4.5-air", torch.bfloat16), (33, "mini", torch.float16),
                                             (64, "glm-4.5", torch.bfloat16), (5, "tiny", torch.float32)])
def test_route_small_bit_identical_to_separate_kernels(T, shape_name, dt):
    lib = _lib.load()
    shape = synth.SHAPES[shape_name]
    gate_w, bias = synth.make_router_weights(shape, dt)
    x = synth.make_tokens(shape, T, dt, 40 + T).cuda()
    gate_w, bias = gate_w.cuda(), bias.cuda()
    E, k, H, tr = shape.E, shape.k, shape.H, 16 if T <= 16 else 32
    cap = lib.glm4moe_tile_capacity(T, k, E, tr)
    code = _lib.DTYPE_CODE[dt]
    e_lo, e_hi, sh_lo, sh_hi = E // 4, E, 0, T   # also exercises the expert-parallel filter

    def bufs():
        return dict(logits=torch.zeros((T, E), dtype=dt, device="cuda"),
                    idx=torch.zeros((T, k), dtype=torch.int32, device="cuda"),
                    w=torch.zeros((T, k), dtype=torch.float32, device="cuda"),
                    off=torch.zeros(E + 1, dtype=torch.int32, device="cuda"),
                    rt=torch.full((T * k,), -7, dtype=torch.int32, device="cuda"),
                    rw=torch.zeros(T * k, dtype=torch.float32, device="cuda"),
                    sr=torch.zeros(T * k, dtype=torch.int32, device="cuda"),
                    tiles=torch.zeros((cap, 4), dtype=torch.int32, device="cuda"),
                    nt=torch.zeros(4, dtype=torch.int32, device="cuda"))
    a, b = bufs(), bufs()
    st = _lib.stream_ptr()
    _lib.check(lib.glm4moe_router_logits(x.data_ptr(), gate_w.data_ptr(), a["logits"].data_ptr(), T, H, E, code, st))
    _lib.check(lib.glm4moe_router_select(a["logits"].data_ptr(), bias.data_ptr(), T, E, k, shape.routed_scaling_factor,
                                         code, a["idx"].data_ptr(), a["w"].data_ptr(), st))
    scr = torch.empty(lib.glm4moe_permute_scratch_elems(T, E), dtype=torch.int32, device="cuda")
    _lib.check(lib.glm4moe_permute(a["idx"].data_ptr(), a["w"].data_ptr(), T, k, E, e_lo, e_hi, tr, sh_lo, sh_hi,
                                   a["off"].data_ptr(), a["rt"].data_ptr(), a["rw"].data_ptr(), a["sr"].data_ptr(),
                                   a["tiles"].data_ptr(), cap, a["nt"].data_ptr(), scr.data_ptr(), st))
    ticket = torch.zeros(40, dtype=torch.int32, device="cuda")
    for _ in range(2):   # twice: the ticket re-arms itself
        _lib.check(lib.glm4moe_route_small(x.data_ptr(), gate_w.data_ptr(), bias.data_ptr(), b["logits"].data_ptr(), T, H,
                                           E, k, shape.routed_scaling_factor, code, e_lo, e_hi, tr, sh_lo, sh_hi,
                                           b["idx"].data_ptr(), b["w"].data_ptr(), b["off"].data_ptr(),
                                           b["rt"].data_ptr(), b["rw"].data_ptr(), b["sr"].data_ptr(),
                                           b["tiles"].data_ptr(), cap, b["nt"].data_ptr(), ticket.data_ptr(), None, None, 0, st))
    torch.cuda.synchronize()
    assert int(ticket.abs().sum()) == 0
    n_rows = int(a["off"][E])
    for name in ("logits", "idx", "w", "off", "sr", "nt"):
        assert torch.equal(a[name], b[name]), name
    assert torch.equal(a["rt"][:n_rows], b["rt"][:n_rows]) and torch.equal(a["rw"][:n_rows], b["rw"][:n_rows])
    nt = int(a["nt"][0])
    assert nt > 0 and torch.equal(a["tiles"][:nt], b["tiles"][:nt])


@pytest.mark.parametrize("shape_name,dt", [("glm-4.5-air", torch.bfloat16), ("mini", torch.float16),
                                           ("glm-4.5", torch.bfloat16)])
def test_merged_front_end_bit_identical_to_two_launches(shape_name, dt):
    """T <= 2: router logits + selection + permutation inside the persistent decode kernel
    (`RouteMerged`, decode_fused.cu: one launch per layer) against route_small_kernel + decode_fused_kernel
    (two launches): logits, indices, weights, permutation, work list and the layer output bit for bit,
    repeatedly (the ticket and the flag re-arm themselves) and under CUDA-graph replay.  GLM-4.5: E = 160 > 148
    CTAs (a CTA computes two experts' logits)."""
    lib = _lib.load()
    shape = synth.SHAPES[shape_name]
    moe = _module(shape, dt, synth.make_weights(shape, dt), max_tokens=64)
    names = ("logits", "idx", "w", "expert_offsets", "slot_row", "num_tiles")
    try:
        for T in (1, 2):
            got = {}
            for merged in (1, 0, 1):
                _lib.check(lib.glm4moe_debug_set(9, merged))
                for rep in range(3):
                    x = synth.make_tokens(shape, T, dt, 700 + 10 * T + rep).cuda()
                    out = moe(x).clone()
                    assert moe.last_launches == (1 if merged else 2), (merged, moe.last_launches)
                    assert moe.last_fused
                    views = {n: moe.workspace_view(n, T).clone() for n in names}
                    n_rows = int(views["expert_offsets"][-1])
                    views["row_token"] = moe.workspace_view("row_token").clone()[:n_rows]
                    views["row_weight"] = moe.workspace_view("row_weight").clone()[:n_rows]
                    views["tiles"] = moe.workspace_view("tiles").clone()[:int(views["num_tiles"].flatten()[0])]
                    key = (T, rep)
                    if merged == 0:
                        ref_out, ref_views = got[key]
                        assert torch.equal(out, ref_out), f"T={T} rep={rep}: outputs differ"
                        for n, v in views.items():
                            assert torch.equal(v, ref_views[n]), (T, rep, n)
                    elif key in got:      # second merged pass: still the same after the un-merged runs in between
                        assert torch.equal(out, got[key][0])
                    got[key] = (out, views)
        # CUDA graph: the merged kernel replays with new tokens
        _lib.check(lib.glm4moe_debug_set(9, 1))
        static_in = synth.make_tokens(shape, 1, dt, 991).cuda()
        moe(static_in)
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            moe(static_in)
        torch.cuda.current_stream().wait_stream(s)
        with torch.cuda.graph(g):
            static_out = moe(static_in)
        for b_ in (992, 993):
            xn = synth.make_tokens(shape, 1, dt, b_).cuda()
            static_in.copy_(xn)
            g.replay()
            torch.cuda.synchronize()
            assert torch.equal(static_out, moe(xn))
    finally:
        lib.glm4moe_debug_set(9, 1)


@pytest.mark.parametrize("shape_name,Ts,fp8", [("mini", (1, 5, 16, 17, 32), False), ("glm-4.5-air", (1, 2, 13, 32), False),
                                               ("mini", (1, 5, 16, 17, 32), True), ("glm-4.5-air", (1, 3, 32), True)])
def test_direct_mode_bit_identical_to_general_path(shape_name, Ts, fp8):
    """T <= 32 runs the fused front-end and feeds the input tokens to the gate_up MMA in place
    (no gathered copy); the general path must give the same bits (same K order per output)."""
    lib = _lib.load()
    shape, dt = synth.SHAPES[shape_name], torch.bfloat16
    moe = _module(shape, dt, synth.make_weights(shape, dt), max_tokens=64)
    modes = {"fused": (1, 1, 2), "direct": (1, 0, 4), "general": (0, 0, 7)}
    if fp8:
        moe.quantize_fp8(drop_native=True)
        modes = {"fused": (1, 1, 2), "direct": (1, 0, 8), "general": (0, 0, 8)}   # no un-fused direct variant
    try:
        for T in Ts:
            x = synth.make_tokens(shape, T, dt, 900 + T).cuda()
            outs = {}
            for name, (direct, fused, launches) in modes.items():
                _lib.check(lib.glm4moe_debug_set(0, direct))
                _lib.check(lib.glm4moe_debug_set(1, fused))
                # T <= 2 (bf16): the front-end runs inside the persistent kernel, ONE launch (test below)
                want = 1 if name == "fused" and T <= 2 and not fp8 else launches
                for rep in range(3):       # repeated: the in-kernel counters re-arm themselves
                    out = moe(x).clone()
                    assert moe.last_launches == want, (name, moe.last_launches)
                    if rep:
                        assert torch.equal(out, outs[name]), f"T={T} {name}: run-to-run difference"
                    outs[name] = out
                outs[name + "_idx"] = moe.workspace_view("idx", T).clone()
            assert torch.equal(outs["fused_idx"], outs["general_idx"])
            assert torch.equal(outs["direct"], outs["general"]), f"T={T}: direct and general paths differ"
            assert torch.equal(outs["fused"], outs["general"]), f"T={T}: fused and general paths differ"
    finally:
        lib.glm4moe_debug_set(0, 1)
        lib.glm4moe_debug_set(1, 1)

# ------------------------------------------------------------------------------------------
# BASELINE configs[1]: GLM-4.5-Air shape, bf16
# ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def air():
    shape, dt = synth.AIR, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    moe = _module(shape, dt, wts, max_tokens=4096)
    return shape, dt, wts, moe


@pytest.mark.parametrize("T", [1, 32, 256])
def test_air_decode_vs_oracle(air, T):
    shape, dt, wts, moe = air
    x = synth.make_tokens(shape, T, dt, T)
    torch.set_num_threads(max(1, torch.get_num_threads()))
    with torch.inference_mode():
        out = moe(x.cuda())
        torch.cuda.synchronize()
        assert moe.last_impl == "umma"
        idx = moe.workspace_view("idx", T).cpu().long()
        w = moe.workspace_view("w", T).cpu().to(dt)
        # routing vs the oracle router on the box's CPU: every differing token must be a boundary pair with a
        # logit rounding flip (_explain_routing), none is excused wholesale
        logits = moe.workspace_view("logits", T).cpu()
        _explain_routing(idx, logits, x, wts["gate_weight"], wts["bias"], shape, dt)
        rt, _, offs = O.permutation(idx, shape.E)
        assert torch.equal(moe.workspace_view("expert_offsets").cpu(), offs)
        assert torch.equal(moe.workspace_view("row_token").cpu()[:T * shape.k], rt)
        truth = O.moe_forward_fp32(x, wts, shape.k, 1.0, routing=(idx, w))["out"]
    nbad, rel = _close(out, truth, RTOL_BF16)
    assert nbad == 0 and rel < 6e-3, (nbad, rel)


def test_air_prefill_properties(air):
    """Full BASELINE size (T=4096): size-independent properties, no CPU oracle needed."""
    shape, dt, wts, moe = air
    T = 4096
    x = synth.make_tokens(shape, T, dt, 77).cuda()
    with torch.inference_mode():
        y = moe(x)
        offs = moe.workspace_view("expert_offsets").cpu().long()
        idx = moe.workspace_view("idx", T).cpu().long()
        rt = moe.workspace_view("row_token").cpu().long()
        sr = moe.workspace_view("slot_row", T).cpu().long()
        assert int(offs[-1]) == T * shape.k
        assert torch.equal(offs[1:] - offs[:-1], torch.bincount(idx.view(-1), minlength=shape.E))
        assert bool((idx[:, 1:] > idx[:, :-1]).all())                 # k distinct experts, ascending
        assert torch.equal(rt[sr.view(-1)], torch.arange(T).repeat_interleave(shape.k))
        for e in (0, 17, 127):                                        # token-ascending inside an expert
            seg = rt[offs[e]:offs[e + 1]]
            assert bool((seg[1:] > seg[:-1]).all())
        assert bool(torch.isfinite(y.float()).all())
        # a decode-size slice of the same tokens (CUDA-core router, 16-row tiles) reproduces the rows
        # of the prefill pass (tensor-core router, 256-row tiles) wherever the routing agrees
        sub = torch.arange(0, T, 128, device="cuda")
        ys = moe(x[sub])
        same = (moe.workspace_view("idx", sub.numel()).cpu().long() == idx[sub.cpu()]).all(-1)
        assert int(same.sum()) >= sub.numel() - 1
        a, b = ys.float().cpu()[same], y[sub].float().cpu()[same]
        assert float((a - b).abs().max()) <= 2e-2 * float(b.abs().max())
        assert float(((a - b).norm() / b.norm())) < 2e-3


@pytest.mark.parametrize("T,knobs,want_rows", [
    (4096, {}, 256),                      # the prefill configuration: 256-token gated tiles + 256x256 down tiles
    (4096, {"down2": 0}, 256),            # plain 128-feature down tiles at 256 tokens
    (4096, {"bk32": 1}, 256),             # K blocks of 32 / 64-byte swizzle ring (opt-in)
    (4000, {}, 256),                      # ragged: T not a multiple of anything
    (1024, {}, 128),                      # mean 64 rows per expert -> 128-token tiles
    (512, {}, 64),                        # mean 32 rows per expert -> 64-token tiles
    (700, {"tile_rows": 128}, 128),       # forced geometry on a batch whose experts mostly hold < 1 tile
    (4096, {"tile_rows": 64}, 64),        # many tiles per expert (weight slab shared by 4+ CTAs)
])
def test_air_prefill_tiles_vs_oracle(air, T, knobs, want_rows):
    """every bf16 tensor-core tile geometry behind the prefill numbers (64 / 128 / 256 token
    rows, gated + `kModeDown2`, the BK=32 ring) against the fp32 oracle under forced routing on sampled
    token rows spread over every expert and over first / last / ragged token tiles."""
    shape, dt, wts, moe = air
    x = synth.make_tokens(shape, T, dt, 1000 + T)
    with _knobs(**knobs), torch.inference_mode():
        out = moe(x.cuda())
        torch.cuda.synchronize()
        assert moe.last_impl == "umma" and not moe.last_fused
        assert _lib.load().glm4moe_layer_tile_rows(moe._handle, T) == want_rows
        idx = moe.workspace_view("idx", T).cpu().long()
        w = moe.workspace_view("w", T).cpu().to(dt)
        logits = moe.workspace_view("logits", T).cpu()
    _explain_routing(idx, logits, x, wts["gate_weight"], wts["bias"], shape, dt)
    rt, _, offs = O.permutation(idx, shape.E)
    assert torch.equal(moe.workspace_view("expert_offsets").cpu(), offs)
    assert torch.equal(moe.workspace_view("row_token").cpu()[:T * shape.k], rt)
    n, rel = _check_sampled_rows(out, x, wts, shape, dt, idx, w, want_rows, seed=T)
    assert n >= 320


@pytest.mark.parametrize("T,knobs", [(4096, {}), (4000, {"down2": 0}), (1024, {}), (700, {"tile_rows": 128}),
                                     (333, {"tile_rows": 256})])
def test_air_prefill_tail_boxes_bit_identical(air, T, knobs):
    """A ragged token tile fetched through a TMA box just large enough for its rows (`TailMaps`,
    umma_gemm.cu) feeds the MMA the same operands as the whole-tile box: outputs bit for bit."""
    shape, dt, wts, moe = air
    x = synth.make_tokens(shape, T, dt, 4400 + T).cuda()
    outs = []
    for tail in (1, 0, 1):
        with _knobs(tail_box=tail, **knobs), torch.inference_mode():
            outs.append(moe(x).clone())
            assert moe.last_impl == "umma" and not moe.last_fused
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    assert bool(torch.isfinite(outs[0].float()).all())


def test_bias_survives_a_dtype_cast_on_the_gpu():
    """build in fp32, `.to(bfloat16).cuda()`: the fp32 bias buffer must not be cast (the library
    reads E * 4 bytes from it); routing against the oracle."""
    shape, dt = synth.MINI, torch.bfloat16
    wts32 = synth.make_weights(shape, torch.float32)
    moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))          # fp32 default dtype
    moe.load_stacked(wts32)
    moe = moe.to(dt).cuda()
    assert moe.gate.e_score_correction_bias.dtype == torch.float32
    x = synth.make_tokens(shape, 48, dt, 3)
    out = moe(x.cuda())
    idx = moe.workspace_view("idx", 48).cpu().long()
    gate_w = wts32["gate_weight"].to(dt)
    oi, _ = O.canonical_routing(*O.router(x, gate_w, wts32["bias"], shape.k, shape.routed_scaling_factor))
    assert torch.equal(idx, oi)
    assert bool(torch.isfinite(out.float()).all())


def test_second_device_after_first():
    """the > 48 KB dynamic shared memory attribute is per device; a layer on cuda:1 after one on
    cuda:0 must launch (needs >= 2 GPUs)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    shape, dt = synth.MINI, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    outs = []
    for T in (8, 200):
        x = synth.make_tokens(shape, T, dt, T)
        per_dev = []
        for d in (0, 1):
            prev = torch.get_default_dtype()
            torch.set_default_dtype(dt)
            try:
                moe = glm4_moe_b200.Glm4MoeMoE(_cfg(shape))
            finally:
                torch.set_default_dtype(prev)
            moe.load_stacked(wts)
            moe = moe.to(f"cuda:{d}")
            per_dev.append(moe(x.to(f"cuda:{d}")).cpu())
        assert torch.equal(per_dev[0], per_dev[1])
        outs.append(per_dev)


# ------------------------------------------------------------------------------------------
# BASELINE configs[2]: FP8 e4m3 weights with 128x128 block scales, per-token 1x128 activations
# ------------------------------------------------------------------------------------------
RTOL_FP8 = 5e-2      # north_star: FP8 outputs within 5e-2 of the fp32 reference (on the dequantised weights)


def _fp8_check(out, x, deq, idx, w, shape, dt, pow2=False):
    """Two gates.  (1) kernel == the quant-tier algorithm (oracle.moe_forward_fp8: identical e4m3
    operands, fp32 math) under the elementwise bf16 criterion, rtol 2e-2.  (2) north_star's FP8
    tolerance against the fp32 reference with *unquantised* activations: e4m3 activations carry
    ~3 % relative rounding noise per GEMM by construction, so 5e-2 is checked on the relative
    Frobenius error and elementwise as |d| <= 5e-2 |ref| + 4 * 5e-2 * rms(ref) (a 4-sigma band of
    that noise, which a Gaussian leaves with probability 6e-5: at most 1e-4 of the elements may;
    the reference torch W8A8 emulation fails the tighter form the same way)."""
    ref_w = {"w_gate_up": deq["w_gate_up"].cpu(), "w_down": deq["w_down"].cpu()}
    algo = O.moe_forward_fp8(x, ref_w, (idx, w), dt, pow2_scales=pow2)
    nbad, rel = _close(out, algo, 2e-2)
    assert nbad == 0 and rel < 6e-3, ("vs quant-tier oracle", nbad, rel)
    truth = O.moe_forward(x.float(), ref_w, shape.k, shape.routed_scaling_factor, routing=(idx, w.float()))["out"]
    got = out.float().cpu()
    rel = float((got - truth).norm() / truth.norm())
    rms = truth.pow(2).mean().sqrt()
    nbad = int(((got - truth).abs() > RTOL_FP8 * truth.abs() + 4 * RTOL_FP8 * rms).sum())
    assert rel < RTOL_FP8 and nbad <= 1e-4 * got.numel(), ("vs fp32 reference", rel, nbad)


def _fp8_case(shape, dt, T, batch, max_tokens=256, pow2=False):
    wts = synth.make_weights(shape, dt)
    moe = _module(shape, dt, wts, max_tokens=max_tokens)
    moe.quantize_fp8(pow2_scales=pow2)
    x = synth.make_tokens(shape, T, dt, batch)
    with torch.inference_mode():
        out = moe(x.cuda())
        torch.cuda.synchronize()
        idx = moe.workspace_view("idx", T).cpu().long()
        w = moe.workspace_view("w", T).cpu().to(dt)
        # routing is untouched by the quant tier (router stays in the model dtype)
        oi, _ = O.canonical_routing(*O.router(x, wts["gate_weight"], wts["bias"], shape.k,
                                              shape.routed_scaling_factor))
        assert torch.equal(idx, oi)
        _fp8_check(out, x, moe.dequantized_weights(), idx, w, shape, dt, pow2)
    return moe


@pytest.mark.parametrize("T", [1, 40, 200, 700])
def test_fp8_block_scaled_mini(T):
    _fp8_case(synth.MINI, torch.bfloat16, T, T, max_tokens=max(T, 256))


@pytest.mark.parametrize("T", [1, 17, 40, 200, 700, 1000])
def test_fp8_pow2_scales_mini(T):
    """Power-of-two (UE8M0) tier: weight block scales and every activation scale are powers of two in all
    regimes (fused decode kernel, swap-AB tiles, prefill); from ~48 rows per expert the GEMMs run with
    hardware block scaling (umma_fp8_mx.cu: tcgen05.mma.kind::mxf8f6f4.block_scale, scales in TMEM)."""
    moe = _fp8_case(synth.MINI, torch.bfloat16, T, T, max_tokens=max(T, 256), pow2=True)
    assert moe._cfg.weight_format == _lib.W_FP8_BLOCK128_POW2


def test_fp8_scale_rounding_error_budget():
    """DESIGN.md section 4 "scale rounding budget": relative Frobenius error of the FP8 layer output against the
    fp32 reference, for the amax/448 tier and the power-of-two tier, (a) on the dequantised weights (north_star's
    5e-2 bar: activation quantisation only) and (b) on the ORIGINAL weights (weight + activation quantisation).
    Rounding the scales up to powers of two may cost at most a quarter on top of the fp32-scale tier.  The
    numbers go to gpurun_out/r2_fp8_error_budget.json when that directory exists."""
    import json
    import os
    shape, dt = synth.MINI, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    ref_w = {"w_gate_up": wts["w_gate_up"].float(), "w_down": wts["w_down"].float()}
    report = {}
    for T in (40, 700):
        x = synth.make_tokens(shape, T, dt, 77 + T)
        for pow2 in (False, True):
            moe = _module(shape, dt, wts, max_tokens=1024)
            moe.quantize_fp8(pow2_scales=pow2)
            with torch.inference_mode():
                out = moe(x.cuda()).float().cpu()
                idx = moe.workspace_view("idx", T).cpu().long()
                w = moe.workspace_view("w", T).cpu().to(dt)
                deq = {k_: v.cpu() for k_, v in moe.dequantized_weights().items()}
                t_deq = O.moe_forward(x.float(), deq, shape.k, shape.routed_scaling_factor,
                                      routing=(idx, w.float()))["out"]
                t_orig = O.moe_forward(x.float(), ref_w, shape.k, shape.routed_scaling_factor,
                                       routing=(idx, w.float()))["out"]
            report[f"T={T} {'pow2' if pow2 else 'amax/448'}"] = {
                "vs_fp32_on_dequantised_weights": float((out - t_deq).norm() / t_deq.norm()),
                "vs_fp32_on_original_weights": float((out - t_orig).norm() / t_orig.norm())}
        a, b = report[f"T={T} amax/448"], report[f"T={T} pow2"]
        for key in a:
            # north_star's 5e-2 is stated against the fp32 reference on the DEQUANTISED weights;
            # the figure on the original weights adds the weight quantisation itself and is reported, bounded at 1e-1
            assert b[key] < (RTOL_FP8 if key == "vs_fp32_on_dequantised_weights" else 1e-1), (T, key, b[key])
            assert b[key] <= 1.25 * a[key] + 1e-3, (T, key, a[key], b[key])
    print(json.dumps(report))
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/r2_fp8_error_budget.json", "w") as f:
            json.dump(report, f, indent=1)


def test_fp8_pow2_tiny_fp16_ragged():
    _fp8_case(synth.TINY, torch.float16, 333, 9, max_tokens=512, pow2=True)


def test_fp8_pow2_format_refuses_other_scales():
    """GLM4MOE_W_FP8_BLOCK128_POW2 is a promise about the checkpoint: amax/448 scales are refused at creation."""
    shape, dt = synth.MINI, torch.bfloat16
    moe = _module(shape, dt, synth.make_weights(shape, dt), max_tokens=64)
    moe.quantize_fp8()
    moe.fp8_pow2_scales = True
    with pytest.raises(_lib.Glm4MoeLibraryError):
        moe(synth.make_tokens(shape, 8, dt, 1).cuda())


@pytest.mark.parametrize("shape_name,T", [("mini", 200), ("mini", 700), ("tiny", 333)])
def test_fp8_hardware_block_scaling_matches_software_scaling(shape_name, T):
    """Same e4m3 operands and the same power-of-two scales through the two prefill kernels: scales applied by
    the tensor pipe (one TMEM accumulator over all of K) vs per-k-block fp32 promotion in the epilogue warps.
    Power-of-two scaling is exact, so the two differ by fp32 summation order only; each is run-to-run
    bit-deterministic; the layer's launch count is the same."""
    lib = _lib.load()
    shape, dt = synth.SHAPES[shape_name], torch.bfloat16
    moe = _module(shape, dt, synth.make_weights(shape, dt), max_tokens=1024)
    moe.quantize_fp8(drop_native=True, pow2_scales=True)
    x = synth.make_tokens(shape, T, dt, 78).cuda()
    try:
        outs, n_launch = {}, {}
        for name, knob in (("pair", 2), ("mx", 1), ("feature_major", 3), ("software", 0)):
            _lib.check(lib.glm4moe_debug_set(8, knob))
            for rep in range(3):
                out = moe(x).clone()
                n_launch[name] = moe.last_launches
                if rep:
                    assert torch.equal(out, outs[name]), f"{name}: run-to-run difference"
                outs[name] = out
        assert n_launch["mx"] == n_launch["software"] == n_launch["pair"] == n_launch["feature_major"], n_launch
        b = outs["software"].float()
        # mx: gate_up token-major (umma_fp8_mx_tm.cu) + down feature-major (umma_fp8_mx.cu); feature_major: both
        # GEMMs feature-major; pair: tcgen05 cta_group::2, 256-token x 256-row tiles (umma_fp8_mx2.cu)
        for name in ("mx", "feature_major", "pair"):
            a = outs[name].float()
            rel = float((a - b).norm() / b.norm())
            nbad, _ = _close(outs[name], b.cpu(), 2e-2)
            assert rel < 2e-3 and nbad == 0, (name, rel, nbad)
    finally:
        lib.glm4moe_debug_set(8, 1)


def test_fp8_block_scaled_tiny_fp16():
    _fp8_case(synth.TINY, torch.float16, 96, 7)


def test_fp8_rejects_ragged_block_shapes():
    # I = 704 is not a multiple of the 128x128 scale block (check_cfg, api.cu)
    moe = _module(synth.SMALL, torch.bfloat16, synth.make_weights(synth.SMALL, torch.bfloat16))
    with pytest.raises(ValueError):
        moe.quantize_fp8()


def test_fp8_quantiser_matches_reference_formula():
    """Weight blocks and the dequantised stacks follow q*scale with scale = amax/448."""
    from glm4_moe_b200 import fp8
    w = torch.randn(3, 256, 384) * 0.02
    q, s = fp8.quantize_block128(w)
    assert q.dtype == torch.float8_e4m3fn and s.shape == (3, 2, 3)
    d = fp8.dequantize_block128(q, s)
    assert float((d - w).abs().max()) <= float(w.abs().max()) * 2 ** -4
    assert torch.allclose(s[1, 0, 2], w[1, :128, 256:].abs().max() / 448)


def test_fp8_air_decode(air):
    shape, dt, wts, _ = air
    moe = _module(shape, dt, wts, max_tokens=64)
    moe.quantize_fp8(drop_native=True)
    for T in (1, 32):
        x = synth.make_tokens(shape, T, dt, 300 + T)
        with torch.inference_mode():
            out = moe(x.cuda())
            torch.cuda.synchronize()
            idx = moe.workspace_view("idx", T).cpu().long()
            w = moe.workspace_view("w", T).cpu().to(dt)
            _fp8_check(out, x, moe.dequantized_weights(), idx, w, shape, dt)


@pytest.mark.parametrize("shape_name,T", [("mini", 200), ("mini", 700), ("tiny", 333)])
def test_fp8_prefill_geometry_matches_swapab_path(shape_name, T):
    """>= 48 rows per expert switches the e4m3 tier to token-major 128 x 256 tiles with the
    activation re-quantiser fused into the gate_up epilogue (umma_fp8_prefill.cu).  Same algorithm
    as the 16..64-token swap-AB kernel: identical e4m3 operands and scales, fp32 promotion per
    128-wide k-block in the same order, so the two paths agree to fp32 rounding of the SiLU input
    (fast-exp) and run-to-run the new path is deterministic."""
    lib = _lib.load()
    shape, dt = synth.SHAPES[shape_name], torch.bfloat16
    moe = _module(shape, dt, synth.make_weights(shape, dt), max_tokens=1024)
    moe.quantize_fp8(drop_native=True)
    x = synth.make_tokens(shape, T, dt, 77).cuda()
    try:
        outs = {}
        n_launch = {}
        for name, knob in (("prefill", 1), ("swapab", 0)):
            _lib.check(lib.glm4moe_debug_set(2, knob))
            for rep in range(3):
                out = moe(x).clone()
                n_launch[name] = moe.last_launches
                if rep:
                    assert torch.equal(out, outs[name]), f"{name}: run-to-run difference"
                outs[name] = out
        # the fused re-quantiser removes the stand-alone activation quantiser launch
        assert n_launch["swapab"] == n_launch["prefill"] + 1, n_launch
        a, b = outs["prefill"].float(), outs["swapab"].float()
        rel = float((a - b).norm() / b.norm())
        nbad, _ = _close(outs["prefill"], b.cpu(), 2e-2)
        assert rel < 2e-3 and nbad == 0, (rel, nbad)
    finally:
        lib.glm4moe_debug_set(2, 1)


class _LazyDequant:
    """Per-expert dequantised fp32 weights on demand (a whole dequantised Air stack is 26 GB)."""

    def __init__(self, q, s):
        self.q, self.s, self.shape = q, s, q.shape

    def __getitem__(self, e):
        from glm4_moe_b200 import fp8
        return fp8.dequantize_block128(self.q[e:e + 1], self.s[e:e + 1])[0].cpu()


def _fp8_sampled_check(moe, out, x, shape, dt, idx, w, tile_rows, seed, pow2=False):
    """The two gates of `_fp8_check` on sampled token rows of a full-size batch."""
    sel = _sample_tokens(idx, shape.E, tile_rows, 48, seed)
    deq = {"w_gate_up": _LazyDequant(moe._wq_gate_up, moe._s_gate_up),
           "w_down": _LazyDequant(moe._wq_down, moe._s_down)}
    xs, ri, rw = x[sel], idx[sel], w[sel]
    got = out[sel.to(out.device)]
    with torch.inference_mode():
        algo = O.moe_forward_fp8(xs, deq, (ri, rw), dt, pow2_scales=pow2)
        truth = O.moe_forward(xs.float(), deq, shape.k, shape.routed_scaling_factor, routing=(ri, rw.float()))["out"]
    nbad, rel = _close(got, algo, 2e-2)
    assert nbad == 0 and rel < 6e-3, ("vs quant-tier oracle", nbad, rel)
    g = got.float().cpu()
    rel = float((g - truth).norm() / truth.norm())
    rms = truth.pow(2).mean().sqrt()
    nbad = int(((g - truth).abs() > RTOL_FP8 * truth.abs() + 4 * RTOL_FP8 * rms).sum())
    assert rel < RTOL_FP8 and nbad <= 1e-4 * g.numel(), ("vs fp32 reference", rel, nbad)
    return int(sel.numel())


@pytest.mark.parametrize("T,pow2", [(4096, False), (1000, False), (4096, True), (1000, True), (4096, 2), (1000, 2)])
def test_fp8_air_prefill_vs_quant_oracle(air, T, pow2):
    """Air shape, FP8 tier at prefill (token-major 128 x 256 tiles, fused re-quantiser): sampled token rows
    against the quant-tier oracle (`moe_forward_fp8`, identical e4m3 operands, fp32 math) at rtol 2e-2 and
    against the fp32 reference on the dequantised weights at north_star's 5e-2 (form: DESIGN.md section 4);
    plus batch-split consistency: rows [0, 1024) of the batch equal a 1024-token forward of those rows to
    fp32 rounding (tile boundaries move, values may not).  `pow2`: the UE8M0 tier, whose prefill GEMMs run with
    hardware block scaling (umma_fp8_mx.cu)."""
    shape, dt, wts, _ = air
    moe = _module(shape, dt, wts, max_tokens=T)
    moe.quantize_fp8(drop_native=True, pow2_scales=bool(pow2))
    x = synth.make_tokens(shape, T, dt, 4242 + T)
    lib = _lib.load()
    try:
        _lib.check(lib.glm4moe_debug_set(8, 2 if pow2 == 2 else 1))      # pow2 == 2: the CTA-pair kernel
        with torch.inference_mode():
            out = moe(x.cuda())
            torch.cuda.synchronize()
            idx = moe.workspace_view("idx", T).cpu().long()
            w = moe.workspace_view("w", T).cpu().to(dt)
            logits = moe.workspace_view("logits", T).cpu()
            n_half = min(1024, T // 2)
            half = moe(x[:n_half].cuda())
    finally:
        lib.glm4moe_debug_set(8, 1)
    _explain_routing(idx, logits, x, wts["gate_weight"], wts["bias"], shape, dt)
    assert torch.isfinite(out.float()).all()
    assert _fp8_sampled_check(moe, out, x, shape, dt, idx, w, 128, T, pow2) >= 320
    d = float((out[:n_half].float() - half.float()).norm() / half.float().norm())
    assert d < 1e-3, d


@pytest.mark.parametrize("shape_name,Ts", [("mini", (1, 7, 32, 200)), ("glm-4.5-air", (3, 32))])
def test_peer_memory_combine_world1_bit_identical(shape_name, Ts):
    """`glm4moe_layer_forward_ep_p2p` with a world of one rank (the workspace is the cudaIpc-exportable
    allocation, every row is local): the expert rows + peer-memory combine kernel must reproduce the
    ordinary forward bit for bit (same order, one rounding), also through the fused decode kernel
    running in rows-only mode (T <= 32), repeatedly (epoch counters re-arm)."""
    import ctypes as C
    shape, dt = synth.SHAPES[shape_name], torch.bfloat16
    wts = synth.make_weights(shape, dt)
    ref = _module(shape, dt, wts, max_tokens=256)
    ep = _module(shape, dt, wts, max_tokens=256, peer_workspace=True)
    handle, off = ep.ipc_handle()
    assert len(handle) == 64
    bases = (C.c_void_p * 1)(ep.workspace_base())
    lib = _lib.load()
    for T in Ts:
        x = synth.make_tokens(shape, T, dt, 50 + T).cuda()
        with torch.inference_mode():
            want = ref(x)
            for rep in range(3):
                got = ep.forward_ep_p2p(x, bases)
                assert torch.equal(got, want), f"T={T} rep={rep}"
            assert torch.equal(ep(x), want)       # the ordinary entry still works on the same workspace
            # decode fast path (T <= 32: rows in epoch-parity buffers, ready flags from the expert kernel, no
            # done round) interleaved with the plain pull form and the sliced form on the same epoch counter
            try:
                for fast, sliced in ((1, -1), (0, -1), (1, -1), (1, -1), (0, 1), (1, 1), (0, 0), (1, -1)):
                    _lib.check(lib.glm4moe_debug_set(7, fast))
                    _lib.check(lib.glm4moe_set_ep_sliced(sliced))
                    assert torch.equal(ep.forward_ep_p2p(x, bases), want), f"T={T} fast={fast} sliced={sliced}"
            finally:
                _lib.check(lib.glm4moe_debug_set(7, 1))
                _lib.check(lib.glm4moe_set_ep_sliced(-1))


# ------------------------------------------------------------------------------------------
# BASELINE configs[4]: GLM-4.5 shape (H=5120, E=160, top-8, I=1536, routed_scaling_factor 2.5)
# ------------------------------------------------------------------------------------------
def test_glm45_shape_decode_and_prefill():
    """One GPU holds the whole bf16 layer (7.6 GB).  Decode T=8 (fused kernel, 40 down feature tiles,
    12 gate_up tiles) against the CPU oracle with forced routing; prefill T=8192 (BASELINE configs[4]:
    256-row tcgen05 tiles, 410 rows per expert) via the size-independent properties and sampled token rows
    against the fp32 oracle; FP8 tier of the same module at both sizes against the quant-tier oracle."""
    shape, dt = synth.GLM45, torch.bfloat16
    wts = synth.make_weights(shape, dt)
    Tp = 8192
    moe = _module(shape, dt, wts, max_tokens=Tp)
    T = 8
    x = synth.make_tokens(shape, T, dt, 8)
    xp = synth.make_tokens(shape, Tp, dt, 9)
    with torch.inference_mode():
        out = moe(x.cuda())
        assert moe.last_fused and moe.last_impl == "umma"
        idx = moe.workspace_view("idx", T).cpu().long()
        w = moe.workspace_view("w", T).cpu().to(dt)
        logits = moe.workspace_view("logits", T).cpu()
        _explain_routing(idx, logits, x, wts["gate_weight"], wts["bias"], shape, dt)
        truth = O.moe_forward_fp32(x, wts, shape.k, shape.routed_scaling_factor, routing=(idx, w))["out"]
        nbad, rel = _close(out, truth, RTOL_BF16)
        assert nbad == 0 and rel < 6e-3, (nbad, rel)
        y = moe(xp.cuda())
        assert _lib.load().glm4moe_layer_tile_rows(moe._handle, Tp) == 256 and not moe.last_fused
        offs = moe.workspace_view("expert_offsets").cpu().long()
        pidx = moe.workspace_view("idx", Tp).cpu().long()
        pw = moe.workspace_view("w", Tp).cpu().to(dt)
        plogits = moe.workspace_view("logits", Tp).cpu()
        assert int(offs[-1]) == Tp * shape.k
        assert torch.equal(offs[1:] - offs[:-1], torch.bincount(pidx.view(-1), minlength=shape.E))
        assert bool(torch.isfinite(y.float()).all())
    _explain_routing(pidx, plogits, xp, wts["gate_weight"], wts["bias"], shape, dt)
    n, _ = _check_sampled_rows(y, xp, wts, shape, dt, pidx, pw, 256, seed=45, n_extra=16)
    assert n >= 320
    # quant tier on the same module: decode against the full quant oracle, prefill on sampled rows
    moe.quantize_fp8(drop_native=True)
    with torch.inference_mode():
        q = moe(x.cuda())
        qi = moe.workspace_view("idx", T).cpu().long()
        qw = moe.workspace_view("w", T).cpu().to(dt)
        assert torch.equal(qi, idx)
        _fp8_sampled_check(moe, q, x, shape, dt, qi, qw, 16, 1)
        qp = moe(xp.cuda())
        torch.cuda.synchronize()
        assert torch.equal(moe.workspace_view("idx", Tp).cpu().long(), pidx)
    _fp8_sampled_check(moe, qp, xp, shape, dt, pidx, pw, 128, 2)


# ------------------------------------------------------------------------------------------
# the step in front of the block: residual add + RMSNorm (layers/layernorm.py)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["rmsnorm_bf16_H4096_T5", "rmsnorm_fp32_H1024_T3", "rmsnorm_fp16_H512_T9"])
def test_rmsnorm_vs_reference_golden(name):
    """`glm4moe_add_rmsnorm` through the RMSNorm mirror against the real reference's outputs.  The residual
    stream (one rounding of an exact fp32 sum) is bit-exact; the normalised output may differ by one
    model-dtype ulp where rsqrtf / the reduction order moves the fp32 scale by an ulp."""
    blob = torch.load(os.path.join(os.path.dirname(__file__), "golden", name + ".pt"))
    dt = blob["x"].dtype
    H = blob["x"].shape[-1]
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dt)
    try:
        norm = glm4_moe_b200.RMSNorm(H, eps=blob["case"]["eps"])
    finally:
        torch.set_default_dtype(prev)
    with torch.no_grad():
        norm.weight.copy_(blob["weight"])
    norm = norm.cuda()
    x, res = blob["x"].cuda(), blob["residual"].cuda()
    with torch.inference_mode():
        y0 = norm(x)
        y1, r1 = norm(x.view(1, *x.shape), res.view(1, *res.shape))      # 3-D input like the runner's
    want_res = blob["x"] + blob["residual"] if dt == torch.float32 else blob["residual_out"]   # fp32: see the oracle test
    assert torch.equal(r1[0].cpu(), want_res)
    for got, want in ((y0, blob["y_plain"]), (y1[0], blob["y_add"])):
        d = (got.float().cpu() - want.float()).abs()
        if dt == torch.float32:      # rsqrtf (2 ulp) + reduction order: a few fp32 ulps on every element
            assert bool((d <= 8 * _ulp(dt) * want.float().abs() + 1e-30).all())
        else:
            assert float((d > 0).float().mean()) < 5e-3
            assert bool((d <= _ulp(dt) * 2 * want.float().abs() + 1e-30).all())
    # rows are independent and the kernel is deterministic
    with torch.inference_mode():
        again, _ = norm(x, res)
    assert torch.equal(again, y1[0])


@pytest.mark.parametrize("shape_name,dt,Ts", [("mini", torch.bfloat16, (1, 2, 7, 32, 33, 300)),
                                              ("mini", torch.float16, (1, 16, 200)),
                                              ("glm-4.5-air", torch.bfloat16, (1, 2, 32, 256))])
def test_forward_prenorm_matches_norm_then_block(shape_name, dt, Ts):
    """`glm4moe_layer_forward_prenorm` (decode_layer.py:98-99 as one call) against the two steps it fuses:
    the norm against the ORACLE (`oracle.rmsnorm`: residual stream bit-exact, normalised tokens within one
    ulp), the block's output BIT-IDENTICAL to `RMSNorm` mirror + `forward` (same kernels, the normalised
    tokens only live in the workspace), in every regime (merged one-launch front end at T <= 2, fused decode
    kernel, general path) and repeatedly."""
    shape = synth.SHAPES[shape_name]
    wts = synth.make_weights(shape, dt)
    moe = _module(shape, dt, wts, max_tokens=512)
    g = torch.Generator().manual_seed(99)
    nw = (1.0 + 0.1 * torch.randn(shape.H, generator=g)).to(dt)
    eps = 1e-5
    for T in Ts:
        x = synth.make_tokens(shape, T, dt, 300 + T)
        res = synth.make_tokens(shape, T, dt, 900 + T)
        want_n, want_r = O.rmsnorm(x, nw, eps, res)
        with torch.inference_mode():
            n_gpu, r_gpu = LN.rmsnorm(x.cuda(), nw.cuda(), eps, res.cuda())
            want_out = moe(n_gpu)
            for rep in range(2):
                out, r_out = moe.forward_prenorm(x.cuda(), res.cuda(), nw.cuda(), eps)
                assert torch.equal(r_out, r_gpu) and torch.equal(out, want_out), f"T={T} rep={rep}"
            # plain RMSNorm form (no residual), 3-D input like the runner's
            out0, r0 = moe.forward_prenorm(x.cuda().view(1, T, -1), None, nw.cuda(), eps)
            assert r0 is None
            assert torch.equal(out0, moe(LN.rmsnorm(x.cuda(), nw.cuda(), eps)))
        assert torch.equal(r_gpu.cpu(), want_r)
        d = (n_gpu.float().cpu() - want_n.float()).abs()
        # two nested roundings (x * rstd, then * weight with |weight| up to ~1.4 here): an inner one-ulp move can
        # come out as two ulps of the result
        assert bool((d <= 4 * _ulp(dt) * want_n.float().abs() + 1e-30).all())
        assert float((d > 0).float().mean()) < 5e-3


def _peer_buffer(nbytes):
    """(ptr, 1 KB-aligned base) of a zeroed cudaIpc-exportable allocation"""
    import ctypes as C
    ptr, hbuf = C.c_void_p(), C.create_string_buffer(64)
    _lib.check(_lib.load().glm4moe_p2p_alloc(nbytes + 1024, C.byref(ptr), hbuf), "p2p_alloc")
    return ptr.value, (ptr.value + 1023) // 1024 * 1024


@pytest.mark.parametrize("H,dt", [(4096, torch.bfloat16), (1024, torch.float16), (5120, torch.bfloat16)])
def test_fused_allreduce_norm_world1(H, dt):
    """`glm4moe_ep_allreduce_add_rmsnorm` with a world of one rank (no peer: the all-reduce is the identity),
    one-shot and two-shot form interleaved on the same epoch counter: bit-identical to `glm4moe_add_rmsnorm`
    (same thread mapping and summation order); weight == NULL returns the input (plain all-reduce).  The
    2/8-rank runs are in test_gpu_ep.py."""
    import ctypes as C
    lib = _lib.load()
    max_tokens = 600
    nbytes = lib.glm4moe_ep_comm_bytes(max_tokens, H, _lib.DTYPE_CODE[dt], 1)
    assert nbytes > 0
    raw, base = _peer_buffer(nbytes)
    bases = (C.c_void_p * 1)(base)
    g = torch.Generator().manual_seed(7)
    nw = (1.0 + 0.1 * torch.randn(H, generator=g)).to(dt).cuda()
    st = _lib.stream_ptr(torch.device("cuda"))
    try:
        for T in (1, 3, 32, 128, 129, 600):
            x = torch.randn(T, H, generator=g).to(dt).cuda()
            res = torch.randn(T, H, generator=g).to(dt).cuda()
            want_n, want_r = LN.rmsnorm(x, nw, 1e-5, res)
            want_p = LN.rmsnorm(x, nw, 1e-5)
            o_n, o_r = O.allreduce_add_rmsnorm([x.cpu()], nw.cpu(), 1e-5, res.cpu())      # the oracle, one rank
            assert torch.equal(o_r, want_r.cpu())
            dn = (want_n.float().cpu() - o_n.float()).abs()
            assert bool((dn <= 4 * _ulp(dt) * o_n.float().abs() + 1e-30).all())
            for mode in (-1, 0, 1, 1, 0, -1):
                _lib.check(lib.glm4moe_debug_set(11, mode))
                out, r_out, plain, copy = (torch.empty_like(x) for _ in range(4))
                args = (bases, 0, 1, max_tokens)
                _lib.check(lib.glm4moe_ep_allreduce_add_rmsnorm(*args, x.data_ptr(), res.data_ptr(), nw.data_ptr(), 1e-5,
                                                                T, H, _lib.DTYPE_CODE[dt], out.data_ptr(),
                                                                r_out.data_ptr(), st))
                _lib.check(lib.glm4moe_ep_allreduce_add_rmsnorm(*args, x.data_ptr(), None, nw.data_ptr(), 1e-5,
                                                                T, H, _lib.DTYPE_CODE[dt], plain.data_ptr(), None, st))
                _lib.check(lib.glm4moe_ep_allreduce_add_rmsnorm(*args, x.data_ptr(), None, None, 0.0,
                                                                T, H, _lib.DTYPE_CODE[dt], copy.data_ptr(), None, st))
                torch.cuda.synchronize()
                assert torch.equal(out, want_n) and torch.equal(r_out, want_r), f"T={T} mode={mode}"
                assert torch.equal(plain, want_p) and torch.equal(copy, x), f"T={T} mode={mode}"
        # argument errors, not crashes
        x = torch.zeros(4, H, dtype=dt, device="cuda")
        assert lib.glm4moe_ep_allreduce_add_rmsnorm(bases, 0, 1, max_tokens, x.data_ptr(), x.data_ptr(), None, 1e-5, 4, H,
                                                    _lib.DTYPE_CODE[dt], x.data_ptr(), x.data_ptr(), st) == -1
        assert lib.glm4moe_ep_allreduce_add_rmsnorm(bases, 0, 1, max_tokens, x.data_ptr(), None, None, 1e-5,
                                                    max_tokens + 1, H, _lib.DTYPE_CODE[dt], x.data_ptr(), None, st) == -1
        assert lib.glm4moe_ep_comm_bytes(16, H, _lib.F32, 1) == 0
    finally:
        _lib.check(lib.glm4moe_debug_set(11, -1))
        torch.cuda.synchronize()
        lib.glm4moe_p2p_free(raw)


def test_forward_ep_p2p_prenorm_world1():
    """`glm4moe_layer_forward_ep_p2p_prenorm` with one rank: fused all-reduce + add + norm kernel, then the
    rows-only block + peer-memory combine; bit-identical to `forward_prenorm` of the ordinary block."""
    import ctypes as C
    shape, dt = synth.SHAPES["mini"], torch.bfloat16
    wts = synth.make_weights(shape, dt)
    ref = _module(shape, dt, wts, max_tokens=256)
    ep = _module(shape, dt, wts, max_tokens=256, peer_workspace=True)
    ep.ipc_handle()
    lib = _lib.load()
    raw, base = _peer_buffer(lib.glm4moe_ep_comm_bytes(256, shape.H, _lib.DTYPE_CODE[dt], 1))

    class Comm:
        bases = (C.c_void_p * 1)(base)
        max_tokens = 256

    bases = (C.c_void_p * 1)(ep.workspace_base())
    nw = (1.0 + 0.05 * torch.randn(shape.H)).to(dt).cuda()
    try:
        for T in (1, 9, 32, 200):
            x = synth.make_tokens(shape, T, dt, 70 + T).cuda()
            res = synth.make_tokens(shape, T, dt, 170 + T).cuda()
            with torch.inference_mode():
                want, want_r = ref.forward_prenorm(x, res, nw, 1e-5)
                for rep in range(3):
                    got, got_r = ep.forward_ep_p2p_prenorm(x, res, nw, 1e-5, bases, Comm)
                    assert torch.equal(got, want) and torch.equal(got_r, want_r), f"T={T} rep={rep}"
    finally:
        torch.cuda.synchronize()
        lib.glm4moe_p2p_free(raw)


# ------------------------------------------------------------------------------------------
# dense MLP through the tensor-core grouped GEMM (SURVEY 8f-2 part ii; decode_layer.py:31-36)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("H,I,dt,Ts", [
    (4096, 10944, torch.bfloat16, (1, 33, 200, 1024)),     # GLM-4.5-Air dense layer 0: I = 85.5 x 128
    (4096, 1368, torch.bfloat16, (7, 300)),                # its tensor-parallel shard at 8 ranks: K tail of 24
    (256, 200, torch.float16, (5, 129)),                   # everything ragged
])
def test_dense_mlp_tensor_cores_vs_oracle(H, I, dt, Ts):
    """`glm4moe_mlp_forward` (one weight pair, all rows, `tiles == nullptr` path of umma_grouped_kernel) against
    the fp32 evaluation of mlp.py:49-54 (oracle.expert_mlp) and against the CUDA-core exact path."""
    from glm4_moe_b200.moe import Glm4MoeDenseMLP, dense_mlp_forward
    g = torch.Generator().manual_seed(H + I)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dt)
    try:
        m = Glm4MoeDenseMLP(H, I, "silu", tp_rank=0, tp_size=1)
    finally:
        torch.set_default_dtype(prev)
    with torch.no_grad():
        m.gate_up_proj.weight.copy_((torch.randn((2 * I, H), generator=g) * 0.02).to(dt))
        m.down_proj.weight.copy_((torch.randn((H, I), generator=g) * 0.02).to(dt))
    m = m.cuda()
    m.impl = "umma"
    for T in Ts:
        x = torch.randn((T, H), generator=g).to(dt)
        with torch.inference_mode():
            y = m(x.cuda())
            y2 = m(x.cuda())
            assert torch.equal(y, y2)                      # deterministic, counters re-armed
            truth = O.expert_mlp(x.float(), m.gate_up_proj.weight.float().cpu(), m.down_proj.weight.float().cpu())
            nbad, rel = _close(y, truth, RTOL_BF16)
            assert nbad == 0 and rel < 6e-3, (T, nbad, rel)
            if T <= 64:
                ys = dense_mlp_forward(x.cuda(), m.gate_up_proj.weight, m.down_proj.weight, impl="simt")
                nbad, rel = _close(y, ys.float().cpu(), RTOL_BF16)
                assert nbad == 0 and rel < 6e-3, ("vs simt", T, nbad, rel)
    # a 3-D input keeps its leading shape (decode_layer.py:100 passes [T, H]; the reference linear layers accept any)
    assert m(torch.zeros((2, 3, H), dtype=dt, device="cuda")).shape == (2, 3, H)
