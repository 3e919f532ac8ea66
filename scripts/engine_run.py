#!/usr/bin/env python
"""BASELINE configs[3]: a random-init GLM4-MoE decoder stack through the reference's OWN engine
(`nanovllm.LLM`: scheduler, block manager, flash-attn attention, CUDA-graph decode runner, sampler) with
this package's sparse block swapped in (`glm4_moe_b200.engine_shim`).  Prints one JSON line:
prefill / decode tokens/s, ms per decode step, the generated token ids (greedy) and a digest of them.

  python scripts/engine_run.py --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32
  python scripts/engine_run.py --preset small --layers 4 --block reference --eager       # A/B host stack

The engine host comes from `baseline/_ref/` (installed by `oracle/make_ref.py`); nothing under it is edited.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import glm4_moe_b200  # noqa: E402,F401
from glm4_moe_b200 import engine_shim as S  # noqa: E402

SMALL = dict(hidden_size=512, n_routed_experts=32, num_experts_per_tok=8, moe_intermediate_size=256,
             intermediate_size=1024, num_attention_heads=8, num_key_value_heads=2, head_dim=64,
             vocab_size=2048, routed_scaling_factor=2.5)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="small", choices=["small", "air"])
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--block", default="ours", choices=["ours", "reference"])
    ap.add_argument("--weights", default="bf16", choices=["bf16", "fp8", "fp8_pow2"])
    ap.add_argument("--eager", action="store_true", help="enforce_eager=True (no CUDA-graph decode)")
    ap.add_argument("--seqs", type=int, default=8)
    ap.add_argument("--prompt-len", type=int, default=24)
    ap.add_argument("--max-tokens", type=int, default=12)
    ap.add_argument("--max-num-seqs", type=int, default=32)
    ap.add_argument("--max-model-len", type=int, default=512)
    ap.add_argument("--max-batched-tokens", type=int, default=2048)
    ap.add_argument("--mem", type=float, default=0.85, help="gpu_memory_utilization (KV cache sizing)")
    ap.add_argument("--model-dir", default="")
    ap.add_argument("--tp", type=int, default=1, help="ranks (tensor_parallel_size): attention / dense MLP "
                    "tensor-parallel as the reference's layers do it, sparse blocks expert-parallel (--ep)")
    ap.add_argument("--ep", default="p2p", choices=["p2p", "allreduce", "none"],
                    help="layout of the sparse blocks when --tp > 1 (none: every rank holds all experts)")
    ap.add_argument("--dense", default="reference", choices=["ours", "reference"],
                    help="dense MLP of layer 0 (decode_layer.py:31-36): ours = Glm4MoeDenseMLP (tensor-core grouped GEMM)")
    ap.add_argument("--fused-norm", action="store_true",
                    help="decoder-layer norms through our kernel; post-attention norm (+ the o_proj all-reduce at "
                         "tp > 1) fused to the block that follows (engine_shim.install(fused_norm=True))")
    args = ap.parse_args()

    nanovllm = S.install(sparse_block=args.block, weight_format=args.weights if args.weights != "bf16" else "native",
                         expert_parallel=None if args.ep == "none" or args.tp == 1 else args.ep, dense_mlp=args.dense,
                         fused_norm=args.fused_norm)
    geom = dict(S.AIR_MODEL if args.preset == "air" else SMALL)
    d = args.model_dir or tempfile.mkdtemp(prefix="glm4moe_model_")
    S.make_model_dir(d, num_hidden_layers=args.layers, max_position_embeddings=max(args.max_model_len, 512), **geom)

    t0 = time.perf_counter()
    llm = nanovllm.LLM(d, enforce_eager=args.eager, tensor_parallel_size=args.tp, max_num_seqs=args.max_num_seqs,
                       max_model_len=args.max_model_len, max_num_batched_tokens=args.max_batched_tokens,
                       gpu_memory_utilization=args.mem)
    build_s = time.perf_counter() - t0
    g = torch.Generator().manual_seed(7)
    V = geom["vocab_size"]
    prompts = [torch.randint(3, V, (args.prompt_len + (i % 5),), generator=g).tolist() for i in range(args.seqs)]
    sp = nanovllm.SamplingParams(temperature=0.0, max_tokens=args.max_tokens, ignore_eos=True)
    for p in prompts:
        llm.add_request(p, sp)
    outputs = {}
    prefill_tok = prefill_s = decode_tok = decode_s = 0.0
    decode_steps = []
    while not llm.is_finished():
        torch.cuda.synchronize()
        t = time.perf_counter()
        out, n = llm.step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t
        if n > 0:
            prefill_tok += n
            prefill_s += dt
        else:
            decode_tok += -n
            decode_s += dt
            decode_steps.append((dt, -n))
        for seq_id, toks in out:
            outputs[seq_id] = toks
    toks = [outputs[k] for k in sorted(outputs)]
    steady = decode_steps[2:] or decode_steps         # the first replays of a graph pay one-time costs
    ms = sorted(s[0] for s in steady)[len(steady) // 2] * 1e3 if steady else None
    mlps = [getattr(l.mlp, "local", l.mlp) for l in llm.model_runner.model.model.layers]
    ours = [m for m in mlps if isinstance(m, glm4_moe_b200.Glm4MoeMoE)]
    line = {"what": "decoder stack through the reference engine (nanovllm.LLM) with the sparse block swapped in",
            "preset": args.preset, "layers": args.layers, "moe_layers": len(mlps) - 1, "block": args.block,
            "weights": args.weights, "cuda_graph_decode": not args.eager, "seqs": args.seqs,
            "dense_mlp": type(llm.model_runner.model.model.layers[0].mlp).__module__.split(".")[0] + "." +
            type(llm.model_runner.model.model.layers[0].mlp).__name__,
            "ranks": args.tp, "parallelism": "single" if args.tp == 1 else
            f"tp{args.tp} attention + dense MLP (reference layers, NCCL all-reduce), sparse blocks " +
            ("replicated" if args.ep == "none" else f"ep{args.tp}-{args.ep}"),
            "experts_per_rank": (ours[0].expert_end - ours[0].expert_begin) if ours else None,
            "prompt_len": args.prompt_len, "max_tokens": args.max_tokens,
            "prefill_tokens_per_s": prefill_tok / prefill_s if prefill_s else None,
            "decode_tokens_per_s": decode_tok / decode_s if decode_s else None,
            "decode_ms_per_step_median": ms, "decode_batch": args.seqs, "build_s": build_s,
            "sparse_blocks_ours": len(ours), "fused_norm": bool(args.fused_norm),
            "fused_post_layers": sum(1 for l in llm.model_runner.model.model.layers if "_fused_post" in l.__dict__),
            "last_impl": ours[0].last_impl if ours else None, "last_fused": ours[0].last_fused if ours else None,
            "gpu_mem_gb": torch.cuda.max_memory_allocated() / 1e9,
            "workspace_allocations": len({m.workspace_base() for m in ours if m._workspace is not None}),
            "digest": hashlib.sha256(json.dumps(toks).encode()).hexdigest()[:16], "token_ids": toks}
    print(json.dumps(line))
    sys.stdout.flush()
    if args.tp > 1:
        try:
            llm.exit()         # tells the spawned ranks to leave their loop (llm_engine.py:37-41)
        except Exception as e:  # noqa: BLE001
            print(f"engine exit: {e}", file=sys.stderr)
    os._exit(0)        # the engine's atexit handler tears NCCL down; nothing left to flush


if __name__ == "__main__":
    main()
