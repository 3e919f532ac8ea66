"""GPU box: the REAL reference module (installed copy, oracle/_ref/, see oracle/make_ref.py) as the checker.

1. the CPU restatement equals the reference bit for bit on THIS machine's host CPU (the pin that
   tests/test_oracle_golden.py runs in the builder container, repeated where the GPU numbers are made);
2. the reference's own `Glm4MoeMoE.forward` (moe.py:128-176) executed on the SAME GPU in the model dtype
   (its `F.linear` -> cuBLAS, `torch.topk`, `index_add_`) against the CUDA path through the C ABI:
   index sets equal except where the two logit computations round a boundary logit differently
   (checked per token), outputs within the reference's own rtol = atol = 1e-2 (nanovllm/tests/moe.py:40)
   and within rtol 2e-2 (+ rtol * rms) of the forced-routing fp32 truth."""
import pytest
import torch

from glm4_moe_b200 import synth
from oracle import moe_oracle as O
from oracle import reference_live as R
from test_oracle_golden import live_pin, FULL
from test_gpu_parity import _module, _close, _ulp, RTOL_BF16

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not R.available(), reason="oracle/_ref missing: run python oracle/make_ref.py")]


@pytest.mark.parametrize("name", FULL)
def test_oracle_bit_exact_vs_live_reference_on_this_box(name):
    live_pin(name)


@pytest.mark.parametrize("shape_name,dt,T", [("mini", torch.bfloat16, 48), ("mini", torch.float16, 300),
                                             ("small", torch.bfloat16, 256), ("tiny", torch.float32, 64)])
def test_cuda_path_vs_real_reference_module_on_the_same_gpu(shape_name, dt, T):
    shape = synth.SHAPES[shape_name]
    wts = synth.make_weights(shape, dt)
    x = synth.make_tokens(shape, T, dt, 11)
    ref = R.build_reference_module(shape, dt, device="cuda")
    R.fill_reference_module(ref, wts)
    moe = _module(shape, dt, wts, max_tokens=512)
    with torch.inference_mode():
        out = moe(x.cuda())
        ridx, rw = ref.gate(x.cuda())
        rout = ref(x.cuda())
        rlogits = torch.nn.functional.linear(x.cuda(), ref.gate.weight)       # moe.py:32
    torch.cuda.synchronize()
    assert rout.shape == out.shape and rout.dtype == out.dtype
    idx = moe.workspace_view("idx", T).cpu().long()
    w = moe.workspace_view("w", T).cpu().to(dt)
    logits = moe.workspace_view("logits", T).cpu().float()
    ri, _ = O.canonical_routing(ridx.cpu().long(), rw.cpu())
    mism = (idx != ri).any(-1)
    dl = (logits - rlogits.cpu().float()).abs()
    sc = logits.sigmoid().to(dt).float() + wts["bias"]
    for t in torch.nonzero(mism).flatten().tolist():
        only = list(set(idx[t].tolist()) ^ set(ri[t].tolist()))
        assert len(only) == 2, f"token {t}: sets differ in more than one expert"
        assert float(dl[t, only].max()) > 0, f"token {t}: routing differs without a logit rounding difference"
        assert abs(float(sc[t, only[0]] - sc[t, only[1]])) <= 2 * _ulp(dt), f"token {t}: not a boundary pair"
    from test_gpu_parity import _log_routing
    _log_routing("layer", shape, dt, T, mism.sum(), "the real reference module on the same GPU")
    assert float(mism.float().mean()) <= 0.02
    same = ~mism
    got, want = out.cpu().float()[same], rout.cpu().float()[same]
    if dt == torch.float32:
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    else:
        torch.testing.assert_close(got, want, rtol=1e-2, atol=1e-2)
    with torch.inference_mode():
        truth = O.moe_forward_fp32(x, wts, shape.k, shape.routed_scaling_factor, routing=(idx, w))["out"]
    nbad, rel = _close(out, truth, 1e-5 if dt == torch.float32 else RTOL_BF16)
    assert nbad == 0, (nbad, rel)
