"""What the fused front end buys (SURVEY 8f-2 part i / 8f-4): decode_layer.py:96-99 from the o_proj output to the
block's output, timed as CUDA-graph replays (the reference engine's decode path replays graphs,
model_runner.py:197-213) of

  ref    dist.all_reduce (N > 1) + the reference's RMSNorm as written (layernorm.py:29-40: ~8 torch launches)
         + the sparse block
  split  dist.all_reduce (N > 1) + glm4moe_add_rmsnorm + the sparse block     (three library calls)
  fused  glm4moe_layer_forward_prenorm (N = 1) / glm4moe_layer_forward_ep_p2p_prenorm (N > 1): one call; the
         all-reduce, residual add and norm are ONE kernel over NVLink peer memory (csrc/ep_allreduce.cu)

and of the front end alone (`front_*`: everything up to the normalised tokens, no block).

  python scripts/front_probe.py                       # 1 GPU
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/front_probe.py

One JSON line per (shape, T) on rank 0.  Times: CUDA events around graph replays, max over ranks."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import glm4_moe_b200  # noqa: E402
from glm4_moe_b200 import synth  # noqa: E402
from glm4_moe_b200 import layernorm as LN  # noqa: E402


def ref_add_rms(x, residual, weight, eps):          # layers/layernorm.py:29-40, op for op
    orig = x.dtype
    x = x.float().add_(residual.float())
    residual = x.to(orig)
    var = x.pow(2).mean(dim=-1, keepdim=True)
    x.mul_(torch.rsqrt(var + eps))
    x = x.to(orig).mul_(weight)
    return x, residual


def time_graph(fn, iters, reps, world):
    """ms per call of `fn`, captured `iters` times in one graph, replayed `reps` times"""
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    side.synchronize()
    if world > 1:
        dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        b.synchronize()
        t.append(a.elapsed_time(b) / iters)
    t.sort()
    ms = torch.tensor([t[len(t) // 2]], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="glm-4.5-air")
    ap.add_argument("--tokens", default="1,8,32,128,1024,4096")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--no-block", action="store_true", help="front end only")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    shape, dt, eps = synth.SHAPES[args.shape], torch.bfloat16, 1e-5
    Ts = [int(t) for t in args.tokens.split(",")]
    H = shape.H
    from transformers.models.glm4_moe import Glm4MoeConfig
    cfg = Glm4MoeConfig(hidden_size=H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                        moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                        routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)
    torch.set_default_dtype(dt)
    comm = ep = moe = None
    with torch.device("cuda"):
        if not args.no_block:
            if world > 1:
                from glm4_moe_b200.distributed import ExpertParallelMoE
                local = glm4_moe_b200.Glm4MoeMoE(cfg, ep_rank=rank, ep_size=world, max_tokens=max(Ts),
                                                 peer_workspace=True)
                local.init_random_(1234)
                ep = ExpertParallelMoE(local, shape.E, layout="p2p")
                ep.setup_p2p(max(Ts))
            else:
                moe = glm4_moe_b200.Glm4MoeMoE(cfg, max_tokens=max(Ts))
                moe.init_random_(1234)
    if world > 1:
        from glm4_moe_b200.distributed import PeerComm
        comm = PeerComm(max(Ts), H, dt)
        if ep is not None:
            ep.comm = comm
    g = torch.Generator(device="cuda").manual_seed(11)
    nw = (1.0 + 0.1 * torch.randn(H, generator=g, device="cuda")).to(dt)
    for T in Ts:
        part = (torch.randn(T, H, generator=g, device="cuda") / world ** 0.5).to(dt)
        res = torch.randn(T, H, generator=g, device="cuda").to(dt)
        block = (lambda x: x) if args.no_block else (ep if world > 1 else moe)

        def f_ref():
            y = part.clone()
            if world > 1:
                dist.all_reduce(y)
            x, r = ref_add_rms(y, res, nw, eps)
            return block(x), r

        def f_split():
            y = part.clone()
            if world > 1:
                dist.all_reduce(y)
            x, r = LN.rmsnorm(y, nw, eps, res)
            return block(x), r

        def f_fused():
            if args.no_block:
                if world > 1:
                    return comm.allreduce_add_rmsnorm(part, res, nw, eps)
                return LN.rmsnorm(part, nw, eps, res)
            if world > 1:
                return ep.forward_prenorm(part, res, nw, eps)
            return moe.forward_prenorm(part, res, nw, eps)

        with torch.inference_mode():
            o_split, r_split = f_split()
            o_fused, r_fused = f_fused()
            same = bool(torch.equal(r_split, r_fused))
            # N = 2: NCCL's a + b is the kernel's sum; beyond that NCCL's order differs, outputs agree to rounding
            d = float((o_split.float() - o_fused.float()).abs().max() / o_split.float().abs().max().clamp_min(1e-9))
            ms = {name: time_graph(fn, args.iters, args.reps, world)
                  for name, fn in (("ref", f_ref), ("split", f_split), ("fused", f_fused))}
        if rank == 0:
            print(json.dumps({"what": "front end + block" if not args.no_block else "front end only",
                              "shape": args.shape, "T": T, "ranks": world, "dtype": "bf16",
                              "ms": {k: round(v, 5) for k, v in ms.items()},
                              "saved_us_vs_ref": round((ms["ref"] - ms["fused"]) * 1e3, 2),
                              "saved_us_vs_split": round((ms["split"] - ms["fused"]) * 1e3, 2),
                              "residual_bit_identical_to_split": same, "out_max_rel_diff_vs_split": d}), flush=True)
    torch.cuda.synchronize()
    if world > 1:
        if ep is not None:
            ep.close_p2p()
        comm.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
