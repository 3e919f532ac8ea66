"""GPU (>= 2 devices): expert-parallel layouts over NCCL against the single-GPU block and the fp32
truth.  Skipped on a 1-GPU box; run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_ep.py -m gpu`."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _collect(procs, q, world, limit_s=600):
    """One result per rank; a rank that dies ends the test at once (the survivors would sit in a collective until
    the limit: on an 8-GPU box that is ten minutes of eight GPUs)."""
    import queue
    import time
    out, t0 = [], time.time()
    while len(out) < world:
        try:
            out.append(q.get(timeout=2))
            continue
        except queue.Empty:
            pass
        dead = [p for p in procs if p.exitcode not in (None, 0)]
        if dead or time.time() - t0 > limit_s:
            for p in procs:
                if p.is_alive():
                    p.kill()
            pytest.fail(f"{len(dead)} of {world} ranks failed (exit codes {[p.exitcode for p in procs]}); "
                        f"{len(out)} finished" if dead else f"no result after {limit_s} s")
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return out



def _worker(rank, world, port, q):
    import glm4_moe_b200
    from glm4_moe_b200 import synth
    from glm4_moe_b200.distributed import ExpertParallelMoE
    from oracle import moe_oracle as O
    from test_gpu_parity import _cfg, _module
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        shape, dt = synth.MINI, torch.bfloat16
        wts = synth.make_weights(shape, dt)
        res = {}
        for T in (1, 40, 300):
            x = synth.make_tokens(shape, T, dt, T)
            single = _module(shape, dt, wts)
            with torch.inference_mode():
                y1 = single(x.cuda())
                idx = single.workspace_view("idx", T).cpu().long()
                w = single.workspace_view("w", T).cpu().to(dt)
                truth = O.moe_forward_fp32(x, wts, shape.k, shape.routed_scaling_factor, routing=(idx, w))["out"]
            for layout in ("allreduce", "alltoall", "p2p"):
                local = _module(shape, dt, wts, ep_rank=rank, ep_size=world, max_tokens=512,
                                peer_workspace=(layout == "p2p"))
                ep = ExpertParallelMoE(local, shape.E, layout=layout)
                with torch.inference_mode():
                    y = ep(x.cuda())
                    if layout == "p2p":
                        # peer-memory combine: same summation order and single rounding as one GPU
                        for rep in range(3):
                            assert torch.equal(ep(x.cuda()), y)
                        assert torch.equal(y, y1), f"T={T}: p2p layout differs from the single-GPU block"
                        # sliced form (each rank reduces its token slice and stores it to the peers),
                        # interleaved with the pull form on the same epoch counters
                        from glm4_moe_b200 import _lib
                        for mode in (1, 0, 1, -1):
                            _lib.check(_lib.load().glm4moe_set_ep_sliced(mode))
                            dist.barrier()
                            ys = ep(x.cuda())
                            assert torch.equal(ys, y1), f"T={T}: sliced mode {mode} differs"
                        # decode fast path (one rendezvous per step) on / off, several steps back to back
                        for fast in (0, 1, 1, 0, 1):
                            _lib.check(_lib.load().glm4moe_debug_set(7, fast))
                            dist.barrier()
                            for rep in range(3):
                                assert torch.equal(ep(x.cuda()), y1), f"T={T}: fast={fast} rep={rep} differs"
                        _lib.check(_lib.load().glm4moe_debug_set(7, 1))
                torch.cuda.synchronize()
                if layout == "p2p":
                    ep.close_p2p()
                got = y.float().cpu()
                rms = truth.pow(2).mean().sqrt()
                bad = int(((got - truth).abs() > 2e-2 * truth.abs() + 2e-2 * rms).sum())
                rel = float((got - truth).norm() / truth.norm())
                d1 = float((got - y1.float().cpu()).abs().max() / y1.float().abs().max())
                res[(T, layout)] = (bad, rel, d1)
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def _worker_front(rank, world, port, q):
    """Fused all-reduce + residual add + RMSNorm over peer memory (csrc/ep_allreduce.cu) and the
    one-call form o_proj partial sums -> norm -> expert-parallel block."""
    import glm4_moe_b200
    from glm4_moe_b200 import synth, _lib
    from glm4_moe_b200 import layernorm as LN
    from glm4_moe_b200.distributed import ExpertParallelMoE, PeerComm
    from oracle import moe_oracle as O
    from test_gpu_parity import _module
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    lib = _lib.load()
    try:
        shape, dt, eps = synth.MINI, torch.bfloat16, 1e-5
        H = shape.H
        comm = PeerComm(3000, H, dt)
        g = torch.Generator().manual_seed(5)
        nw = (1.0 + 0.1 * torch.randn(H, generator=g)).to(dt).cuda()
        checked = 0

        def parts_of(T, seed):
            gg = torch.Generator().manual_seed(seed)
            parts = [(0.5 * torch.randn(T, H, generator=gg)).to(dt) for _ in range(world)]
            res = torch.randn(T, H, generator=gg).to(dt).cuda()
            acc = torch.zeros(T, H)
            for p_ in parts:                     # fp32, ascending rank: the kernel's order
                acc += p_.float()
            return parts, res, acc.to(dt).cuda()

        # T < world, ragged slices, the one-shot capacity edge (1 MiB = 1024 rows of 512 bf16) and beyond it
        for T in (1, 3, 32, 130, 1024, 1025, 3000):
            parts, res, y = parts_of(T, 1000 + T)
            mine = parts[rank].cuda()
            want_n, want_r = LN.rmsnorm(y, nw, eps, res)
            want_p = LN.rmsnorm(y, nw, eps)
            # the oracle's restatement of linear.py:153-154 + layernorm.py:29-40 on the CPU: the all-reduced sum and
            # the residual stream bit for bit, the normalised tokens within the two nested roundings
            o_n, o_r = O.allreduce_add_rmsnorm(parts, nw.cpu(), eps, res.cpu())
            assert torch.equal(O.allreduce_add_rmsnorm(parts, None, eps), y.cpu()) and torch.equal(o_r, want_r.cpu())
            dn = (want_n.float().cpu() - o_n.float()).abs()
            assert bool((dn <= 2.0 ** -6 * o_n.float().abs() + 1e-30).all()) and float((dn > 0).float().mean()) < 5e-3
            for mode in (-1, 0, 1, 1, 0, -1):
                _lib.check(lib.glm4moe_debug_set(11, mode))
                dist.barrier()
                for rep in range(2):
                    out, r_out = comm.allreduce_add_rmsnorm(mine, res, nw, eps)
                    assert torch.equal(out, want_n) and torch.equal(r_out, want_r), f"T={T} mode={mode} rep={rep}"
                assert torch.equal(comm.allreduce_add_rmsnorm(mine, None, nw, eps), want_p), f"T={T} mode={mode}"
                assert torch.equal(comm.all_reduce(mine), y), f"T={T} mode={mode}: plain all-reduce"
                checked += 4
            if world == 2:        # a + b in fp32, one rounding: what NCCL computes for two ranks
                z = mine.clone()
                dist.all_reduce(z)
                assert torch.equal(z, y)
        _lib.check(lib.glm4moe_debug_set(11, -1))
        # CUDA-graph capture + replay (epoch and tickets live in the comm buffer)
        parts, res, y = parts_of(32, 77)
        mine = parts[rank].cuda()
        want_n, want_r = LN.rmsnorm(y, nw, eps, res)
        dist.barrier()
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            comm.allreduce_add_rmsnorm(mine, res, nw, eps)
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            g_out, g_res = comm.allreduce_add_rmsnorm(mine, res, nw, eps)
        for rep in range(3):
            g_out.zero_()
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(g_out, want_n) and torch.equal(g_res, want_r), f"graph replay {rep}"
        # o_proj partial sums -> fused all-reduce + add + norm -> expert-parallel block, one library call
        wts = synth.make_weights(shape, dt)
        single = _module(shape, dt, wts, max_tokens=512)
        local = _module(shape, dt, wts, ep_rank=rank, ep_size=world, max_tokens=512, peer_workspace=True)
        ep = ExpertParallelMoE(local, shape.E, layout="p2p")
        ep.comm = comm
        local_nccl = _module(shape, dt, wts, ep_rank=rank, ep_size=world, max_tokens=512)
        ep_nccl = ExpertParallelMoE(local_nccl, shape.E, layout="allreduce")
        for T in (1, 5, 32, 300):
            parts, res, y = parts_of(T, 2000 + T)
            mine = parts[rank].cuda()
            with torch.inference_mode():
                want, want_r = single.forward_prenorm(y, res, nw, eps)
                for rep in range(3):
                    got, got_r = ep.forward_prenorm(mine, res, nw, eps)
                    assert torch.equal(got_r, want_r), f"T={T} rep={rep}: residual stream"
                    assert torch.equal(got, want), f"T={T} rep={rep}: block output differs from one GPU"
                got2, got2_r = ep_nccl.forward_prenorm(mine, res, nw, eps)     # the same steps over NCCL
                if world == 2:
                    assert torch.equal(got2_r, want_r)
                # Beyond two ranks NCCL sums the bf16 partials in another order with other roundings: the residual
                # stream agrees to an ulp with the peer-memory kernel's.  The block's OUTPUT is therefore compared on
                # the input this path itself normalised (NCCL's sum, the same norm kernel), where no token can route
                # differently: the single-GPU block on that input, rows within the expert-parallel all-reduce's
                # own rounding (its partial outputs are rounded per rank, then summed), every row, no exemptions.
                # (NCCL rounds its running sums to bf16: up to world - 1 roundings, each at most half an ulp, 2^-9, of
                # a running sum no larger than sum_r |partial_r|; then the residual add rounds once more.  Measured on
                # 8 ranks: 2^-5 on elements whose partials cancel.)
                mag = sum(p_.float().abs() for p_ in parts).cuda()
                dr = (got2_r.float() - want_r.float()).abs()
                bound = world * 2.0 ** -9 * mag + 2.0 ** -7 * want_r.float().abs() + 2.0 ** -10
                assert bool((dr <= bound).all()), (float(dr.max()), float((dr - bound).max()))
                y2 = mine.clone()
                dist.all_reduce(y2)
                x2, r2 = LN.rmsnorm(y2, nw, eps, res)
                got3 = ep_nccl(x2)
                if torch.equal(r2, got2_r):          # NCCL chose the same sum order twice: the one-call form agrees
                    assert torch.equal(got3, got2)
                want2 = single(x2)
                idx2 = single.workspace_view("idx", T).clone()
                assert torch.equal(local_nccl.workspace_view("idx", T), idx2), f"T={T}: routing on identical input"
                got2 = got3
                row_d = (got2.float() - want2.float()).abs().amax(1) / want2.float().abs().max()
                assert float(row_d.max()) < 3e-2, (T, row_d.tolist()[:8])
                # and that output against the fp32 truth under the routing the kernel chose (north_star rtol 2e-2)
                truth = O.moe_forward_fp32(x2.cpu(), wts, shape.k, shape.routed_scaling_factor,
                                           routing=(idx2.cpu().long(),
                                                    single.workspace_view("w", T).cpu().to(dt)))["out"]
                rms = truth.pow(2).mean().sqrt()
                g2 = got2.float().cpu()
                assert int(((g2 - truth).abs() > 2e-2 * truth.abs() + 2e-2 * rms).sum()) == 0
            checked += 4
        torch.cuda.synchronize()
        ep.close_p2p()
        comm.close()
        q.put((rank, checked))
    finally:
        dist.destroy_process_group()


def test_fused_allreduce_norm_and_prenorm_block_two_gpus():
    """world = 2 by default; GLM4MOE_TEST_WORLD=4/8 for more ranks of one box."""
    world = int(os.environ.get("GLM4MOE_TEST_WORLD", "2"))
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_front, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = _collect(procs, q, world)
    assert all(n > 0 for _, n in out)


def test_expert_parallel_two_gpus():
    """world = 2 by default; GLM4MOE_TEST_WORLD=4/8 runs the same checks over more ranks of one box."""
    world = int(os.environ.get("GLM4MOE_TEST_WORLD", "2"))
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = _collect(procs, q, world)
    for rank, res in out:
        for key, (bad, rel, d1) in res.items():
            assert bad == 0 and rel < 8e-3, (rank, key, bad, rel)
            assert d1 < 3e-2, (rank, key, d1)


def _worker_shared(rank, world, port, q):
    """Several expert-parallel blocks on ONE peer-mapped workspace (`Glm4MoeMoE.share_workspace`, what the engine
    shim uses for the 45 sparse layers of a model): called one after the other like the layers of a decoder stack,
    each must stay bit-identical to its single-GPU block, step after step, in every combine mode."""
    import glm4_moe_b200
    from glm4_moe_b200 import synth, _lib
    from glm4_moe_b200.distributed import ExpertParallelMoE
    from test_gpu_parity import _module
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    lib = _lib.load()
    try:
        shape, dt = synth.MINI, torch.bfloat16
        n_layers = 3
        wts = [synth.make_weights(shape, dt, seed=4000 + i) for i in range(n_layers)]
        singles = [_module(shape, dt, w, max_tokens=512) for w in wts]
        glm4_moe_b200.Glm4MoeMoE.share_workspace = True
        try:
            eps_ = []
            for w in wts:
                local = _module(shape, dt, w, ep_rank=rank, ep_size=world, max_tokens=512, peer_workspace=True)
                ep = ExpertParallelMoE(local, shape.E, layout="p2p")
                ep.setup_p2p(512)
                eps_.append(ep)
            assert len({e.local._ipc_ptr for e in eps_}) == 1 and len(ExpertParallelMoE._p2p_maps) == 1
            assert len({e.local.workspace_base() for e in eps_}) == 1
            checked = 0
            for T in (1, 7, 32, 40, 300, 512):
                x = synth.make_tokens(shape, T, dt, T).cuda()
                with torch.inference_mode():
                    want = []
                    h = x
                    for sgl in singles:                      # a small "decoder stack": each block feeds the next
                        h = sgl(h)
                        want.append(h.clone())
                    for mode, fast in ((-1, 1), (1, 1), (0, 0), (-1, 1)):
                        _lib.check(lib.glm4moe_set_ep_sliced(mode))
                        _lib.check(lib.glm4moe_debug_set(7, fast))
                        dist.barrier()
                        for step in range(3):
                            h = x
                            for li, ep in enumerate(eps_):
                                h = ep(h)
                                assert torch.equal(h, want[li]), f"T={T} mode={mode} fast={fast} step={step} layer={li}"
                                checked += 1
                _lib.check(lib.glm4moe_set_ep_sliced(-1))
                _lib.check(lib.glm4moe_debug_set(7, 1))
            # a block that joins after forwards ran would reset the cross-GPU counters: refused, on every rank alike
            late = _module(shape, dt, wts[0], ep_rank=rank, ep_size=world, max_tokens=512, peer_workspace=True)
            with pytest.raises(_lib.Glm4MoeLibraryError):
                late.ipc_handle(512)
            torch.cuda.synchronize()
            for ep in eps_:
                ep.close_p2p()
            assert not ExpertParallelMoE._p2p_maps
        finally:
            glm4_moe_b200.Glm4MoeMoE.share_workspace = False
        q.put((rank, checked))
    finally:
        dist.destroy_process_group()


def test_expert_parallel_blocks_share_one_peer_workspace():
    """world = 2 by default; GLM4MOE_TEST_WORLD=4/8 for more ranks of one box."""
    world = int(os.environ.get("GLM4MOE_TEST_WORLD", "2"))
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_shared, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = _collect(procs, q, world)
    assert all(n > 0 for _, n in out)


_TIMEOUT_SCRIPT = r'''
import ctypes as C, json, sys
import torch
sys.path.insert(0, %(root)r)
import glm4_moe_b200
from glm4_moe_b200 import _lib
lib = _lib.load()
T, H, world = 4, 512, 2
n = lib.glm4moe_ep_comm_bytes(64, H, _lib.BF16, world)
bufs = []
for _ in range(world):                    # "rank 1" is a second local buffer nobody ever writes flags into
    p, h = C.c_void_p(), C.create_string_buffer(64)
    _lib.check(lib.glm4moe_p2p_alloc(n + 1024, C.byref(p), h))
    bufs.append((p.value + 1023) // 1024 * 1024)
assert _lib.ep_last_error() is None
_lib.check(lib.glm4moe_ep_set_timeout_ms(%(ms)d))
x = torch.randn(T, H, device="cuda").to(torch.bfloat16)
out = torch.empty_like(x)
bases = (C.c_void_p * world)(*bufs)
rc = lib.glm4moe_ep_allreduce_add_rmsnorm(bases, 0, world, 64, x.data_ptr(), None, None, 1e-5, T, H, _lib.BF16,
                                          out.data_ptr(), None, _lib.stream_ptr())
res = {"launch_rc": rc}
try:
    torch.cuda.synchronize()
    res["sync"] = "ok"
except RuntimeError as e:
    res["sync"] = str(e)[:200]
res["ep"] = _lib.ep_last_error()          # no CUDA call: works on the poisoned context
try:
    _lib.check(719, "after the trap")
except _lib.Glm4MoeLibraryError as e:
    res["message"] = str(e)
print(json.dumps(res))
sys.stdout.flush()
import os
os._exit(0)
'''


def test_a_wait_that_gives_up_is_reported_through_the_host_record():
    """One GPU: an all-reduce over two 'ranks' of which the second never shows up.  With the limit lowered to 50 ms
    the kernel gives up instead of spinning for 18 s; the trap that follows kills the CUDA context of that process
    (hence the subprocess), and `glm4moe_ep_last_error` still says which rank waited for whom."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", _TIMEOUT_SCRIPT % {"root": root, "ms": 50}], capture_output=True,
                       text=True, timeout=300)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, (r.stdout[-1000:], r.stderr[-2000:])
    res = json.loads(lines[-1])
    assert res["launch_rc"] == 0 and res["sync"] != "ok", res
    ep = res["ep"]
    assert ep is not None and ep["code"] == -5 and ep["rank"] == 0 and ep["peer"] == 1, res
    assert ep["site"] == "all-reduce + norm" and ep["waited_for_call"] == 1 and ep["last_flag"] == 0, res
    assert "timed out" in res["message"] and "rank 0 waited for call 1 of rank 1" in res["message"], res
