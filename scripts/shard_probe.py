"""One GPU standing in for ONE rank of an expert-parallel group: `Glm4MoeMoE(ep_rank=r, ep_size=R).forward`
(local experts + the shared expert on the rank's token slice, partial output) at decode sizes, with the
speculative expert prefetch (GLM4MOE_KNOB_SPEC_PREFETCH) off / auto / n.  The cross-rank combine is not in it:
this isolates what a rank's own kernels cost per step.

  python scripts/shard_probe.py [--ranks 8] [--tokens 32] [--spec -1,0,1,2,3]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import glm4_moe_b200  # noqa: E402
from glm4_moe_b200 import synth, _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="glm-4.5-air")
    ap.add_argument("--ranks", default="8,1")
    ap.add_argument("--tokens", default="32,16")
    ap.add_argument("--spec", default="0,-1,1,2,3")
    ap.add_argument("--weights", default="bf16", choices=["bf16", "fp8"])
    ap.add_argument("--steps", type=int, default=200)
    args = ap.parse_args()
    shape, dt = synth.SHAPES[args.shape], torch.bfloat16
    from transformers.models.glm4_moe import Glm4MoeConfig
    cfg = Glm4MoeConfig(hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                        moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                        routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)
    torch.set_default_dtype(dt)
    lib = _lib.load()
    for R in (int(r) for r in args.ranks.split(",")):
        with torch.device("cuda"):
            moe = glm4_moe_b200.Glm4MoeMoE(cfg, ep_rank=0, ep_size=R, max_tokens=64)
            moe.init_random_(1234)
        if args.weights == "fp8":
            moe.quantize_fp8(drop_native=True)
        for T in (int(t) for t in args.tokens.split(",")):
            xs = [synth.make_tokens(shape, T, dt, 4321 + b).cuda() for b in range(16)]
            for spec in (int(s) for s in args.spec.split(",")):
                _lib.check(lib.glm4moe_debug_set(12, spec))
                with torch.inference_mode():
                    side = torch.cuda.Stream()
                    with torch.cuda.stream(side):
                        for i in range(20):
                            moe(xs[i % 16])
                    side.synchronize()
                    graph = torch.cuda.CUDAGraph()         # 16 steps (one per batch) per replay: no host time in it
                    with torch.cuda.graph(graph):
                        for b in range(16):
                            moe(xs[b])
                    for _ in range(3):
                        graph.replay()
                    torch.cuda.synchronize()
                    reps = max(args.steps // 16, 5)
                    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
                    ev[0].record()
                    for i in range(reps):
                        graph.replay()
                        ev[i + 1].record()
                    torch.cuda.synchronize()
                t = sorted(ev[i].elapsed_time(ev[i + 1]) / 16 for i in range(reps))
                print(json.dumps({"what": "one expert-parallel rank's local step" if R > 1 else "single-GPU block",
                                  "shape": args.shape, "weights": args.weights, "ranks": R,
                                  "local_experts": shape.E // R, "T": T, "spec_prefetch": spec,
                                  "ms_median": round(t[len(t) // 2], 5), "ms_p10": round(t[len(t) // 10], 5),
                                  "ms_mean": round(sum(t) / len(t), 5)}), flush=True)
        del moe
        torch.cuda.empty_cache()
    _lib.check(lib.glm4moe_debug_set(12, -1))


if __name__ == "__main__":
    main()
