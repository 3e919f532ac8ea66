#!/usr/bin/env python
"""bench.py — GLM-4.5-Air MoE-layer tokens/s on B200 (BASELINE.json metric, configs[1]).

A "step" is one forward of the sparse-expert block over one batch of T synthetic tokens
(default T=32, the top of north_star's decode 1-32 range).  One JSON line on stdout:

  value      tokens/s with the token batches already resident in HBM (CUDA events, K steps)
  e2e        same metric through the module's host-buffer entry (pinned host in, pinned host
             out, H2D + D2H inside the timed region)
  roofline   HBM roofline of the dominant kernel (gate_up grouped GEMM), timed live with CUDA
             events between stages (glm4moe_layer_forward_timed)
  cpu_baseline  the oracle port of the reference's torch path on the box's host cores (N=1, rank 0)

`--impl reference` times the reference's CPU torch path (oracle port — the reference itself
cannot travel to the GPU box) on the same workload.  N>1: expert-parallel (experts sharded over
ranks, tokens replicated; the weighted expert rows are combined and reduced across ranks by one kernel
reading NVLink peer memory, `--ep-layout allreduce` = NCCL instead), strong scaling.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

import glm4_moe_b200  # noqa: E402,F401
from glm4_moe_b200 import synth  # noqa: E402

METRIC = "GLM-4.5-Air MoE-layer tokens/s"
UNIT = "tokens/s"
N_BATCHES = 16          # distinct token batches rotated through the timed region


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], "measured (MEASURED_PEAKS.json)", d.get("bf16_tflops", 1650.0)
    return 6650.0, "fallback (B200_PROFILING.md)", 1650.0


class ClockSampler:
    """Samples SM clocks / throttle reasons during the timed region (nvidia-smi query fields)."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons = index, [], set()
        self.max_mhz, self._stop, self._thr = None, threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.02)

    def __enter__(self):
        if self._nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def _hf_config(shape):
    from transformers.models.glm4_moe import Glm4MoeConfig
    return Glm4MoeConfig(hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                         moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                         routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)


def _cpu_reference_run(shape, dt, weights, T, steps, warmup):
    """The reference's CPU torch path (oracle port) on the host cores. Returns (tok/s, ms/step)."""
    from oracle import moe_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    xs = [synth.make_tokens(shape, T, dt, b) for b in range(min(N_BATCHES, steps + warmup))]
    with torch.inference_mode():
        for i in range(warmup):
            O.moe_forward(xs[i % len(xs)], weights, shape.k, shape.routed_scaling_factor)
        t0 = time.perf_counter()
        for i in range(steps):
            O.moe_forward(xs[(warmup + i) % len(xs)], weights, shape.k, shape.routed_scaling_factor)
        dt_s = time.perf_counter() - t0
    return T * steps / dt_s, dt_s / steps * 1e3, cores


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--tokens", type=int, default=32, help="tokens per step (decode batch)")
    ap.add_argument("--shape", default="glm-4.5-air")
    ap.add_argument("--sweep", action="store_true", help="also report T in {1..256, 4096} in `sweep`")
    ap.add_argument("--cpu-steps", type=int, default=0, help="cpu_baseline steps (0 = about 15 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ep-layout", default="p2p", choices=["p2p", "allreduce", "alltoall"],
                    help="expert-parallel layout for --gpus > 1 (distributed.py): p2p = combine + reduction in "
                         "one kernel over NVLink peer memory, allreduce / alltoall = NCCL")
    ap.add_argument("--weights", default="bf16", choices=["bf16", "fp8"],
                    help="fp8 = e4m3 expert weights with 128x128 fp32 block scales, per-token 1x128 activations "
                         "(BASELINE configs[2]); single GPU")
    ap.add_argument("--ep-sliced", type=int, default=-1, choices=[-1, 0, 1],
                    help="p2p layout: -1 pick by size, 0 every rank pulls every row, 1 reduce the own token slice "
                         "and store it to the peers (glm4moe_set_ep_sliced)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    shape = synth.SHAPES[args.shape]
    dt = torch.bfloat16
    T = args.tokens
    regime = "decode batch" if T <= 256 else "prefill"
    workload = (f"{shape.name} MoE layer (H={shape.H}, E={shape.E}, top-{shape.k}, I={shape.I}, 1 shared) "
                f"{'bf16' if args.weights == 'bf16' else 'FP8 e4m3 block-128 expert weights, bf16 activations in/out,'} "
                f"random-init, {regime} T={T}")
    fp8 = args.weights == "fp8"
    if fp8 and world > 1:
        raise SystemExit("--weights fp8 is a single-GPU bench")
    n_batches = N_BATCHES if T <= 1024 else 4

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return
        weights = synth.make_weights(shape, dt)
        steps = max(1, min(args.steps, 8))
        warm = max(1, min(args.warmup, 2))
        tps, ms, cores = _cpu_reference_run(shape, dt, weights, T, steps, warm)
        line = {"metric": METRIC, "value": tps, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
                "steps": steps, "warmup": warm, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": workload, "tokens_per_step": T},
                "cpu_baseline": {"value": tps, "unit": UNIT, "cores": cores, "kind": "port",
                                 "sample": f"{steps} steps of the full workload after {warm} warm-up "
                                           "(oracle port of moe.py:128-176; steps capped at 8)"},
                "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    # ------------------------------------------------------------------ our arm (GPU)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the GLM4-MoE path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from __graft_entry__ import build
    if rank == 0:
        build()
    if world > 1:
        dist.barrier()

    torch.set_default_dtype(dt)
    torch.set_default_device(dev)
    moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=max(T, 256), ep_rank=rank, ep_size=world,
                                   peer_workspace=(world > 1 and args.ep_layout == "p2p"))
    torch.set_default_device("cpu")
    torch.set_default_dtype(torch.float32)
    weights = synth.make_weights_device(shape, dt, dev, experts=range(moe.expert_begin, moe.expert_end))
    moe.load_local_stacks(weights)
    if fp8:
        moe.quantize_fp8(drop_native=True)      # offline conversion (fp8.quantize_block128), outside the timed region
    if world > 1:
        from glm4_moe_b200.distributed import ExpertParallelMoE
        layer = ExpertParallelMoE(moe, shape.E, layout=args.ep_layout)
        if args.ep_layout == "p2p":
            from glm4_moe_b200 import _lib
            _lib.check(_lib.load().glm4moe_set_ep_sliced(args.ep_sliced))
            layer.setup_p2p()
    else:
        layer = moe
    batches_host = [synth.make_tokens(shape, T, dt, b).pin_memory() for b in range(n_batches)]
    batches = [b.to(dev) for b in batches_host]

    def step(x):
        return layer(x)

    with torch.inference_mode():
        for i in range(max(args.warmup, 3)):
            step(batches[i % n_batches])
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(local_rank) as clk:
            torch.cuda.synchronize()
            start.record()
            for i in range(args.steps):
                step(batches[i % n_batches])
            end.record()
            torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms_total = start.elapsed_time(end)
        if world > 1:
            t = torch.tensor([ms_total], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_total = float(t.item())
        ms_step = ms_total / args.steps
        launches = moe.last_launches * args.steps
        fused, impl_name = moe.last_fused, moe.last_impl     # of the timed workload (the sweep below runs other T)

        # per-stage timing (separate loop; an event between stages)
        stage_acc = {}
        n_timed = min(args.steps, 50)
        for i in range(n_timed):
            _, ms = moe.forward_timed(batches[i % n_batches])
            for k_, v in ms.items():
                stage_acc[k_] = stage_acc.get(k_, 0.0) + v / n_timed
        # experts hit per batch (for the algorithmic bytes)
        hits = []
        for b in batches:
            moe(b)
            idx = moe.workspace_view("idx", T)
            local = idx[(idx >= moe.expert_begin) & (idx < moe.expert_end)]
            hits.append(int(torch.unique(local).numel()))
        U = sum(hits) / len(hits)

        # end-to-end through the public API with host buffers: pinned host tokens in, result on the host
        out_host = torch.empty((T, shape.H), dtype=dt).pin_memory()

        def e2e_step(i):
            if world == 1:
                moe.forward_host(batches_host[i % n_batches], out_host)     # glm4moe_layer_forward_host
            else:
                y = layer(batches_host[i % n_batches].to(dev, non_blocking=True))
                if rank == 0:
                    out_host.copy_(y, non_blocking=True)
                torch.cuda.synchronize()

        for i in range(3):
            e2e_step(i)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(i)
        e2e_s = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([e2e_s], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s = float(t.item())
        e2e = {"value": T * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": T * shape.H * 2 * world,
               "d2h_bytes_per_step": T * shape.H * 2, "ms_per_step": e2e_s / args.steps * 1e3}

        sweep = None
        if args.sweep and world == 1:
            sweep = []
            for Ts in (1, 2, 4, 8, 16, 32, 64, 128, 256, 4096):
                xs = [synth.make_tokens(shape, Ts, dt, 100 + b).to(dev) for b in range(8)]
                for i in range(5):
                    moe(xs[i % 8])
                torch.cuda.synchronize()
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                n = 100 if Ts <= 256 else 20
                s0.record()
                for i in range(n):
                    moe(xs[i % 8])
                s1.record()
                torch.cuda.synchronize()
                ms = s0.elapsed_time(s1) / n
                us = []
                for x in xs:
                    moe(x)
                    us.append(int(torch.unique(moe.workspace_view("idx", Ts)).numel()))
                Us = sum(us) / len(us)
                by = shape.algorithmic_bytes(Ts, Us)
                fl = shape.flops_per_token() * Ts
                sweep.append({"T": Ts, "ms": ms, "tokens_per_s": Ts / ms * 1e3, "experts_hit": Us,
                              "GBps": by / ms / 1e6, "TFLOPs": fl / ms / 1e9,
                              "tile_rows": moe.last_impl})

    if world > 1 and args.ep_layout == "p2p":
        layer.close_p2p()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, peak_src, tc_peak = _peaks()
    if fp8:
        tc_peak *= 2.0      # no measured FP8 figure in MEASURED_PEAKS.json: dense e4m3 = 2 x the bf16 rate
    H, I, k = shape.H, shape.I, shape.k
    # dominant kernel: gate_up grouped GEMM. Algorithmic bytes of one launch = gate_up weights of
    # the (U local + shared) experts it touches + the activation rows it reads + act rows it writes.
    n_sh = 1
    wb = 1 if fp8 else 2          # bytes per expert-weight element (and per gathered / activation row element)
    rows = T * k / world + T / world
    gu_ms = stage_acc["gate_up"]
    if fused:
        # decode regime: gate_up + SiLU*mul + down + combine are ONE persistent launch
        # (decode_fused.cu); the "gate_up" stage interval brackets exactly that launch.
        kern = "decode_fused_kernel<gate_up+silu_mul+down+combine>"
        gu_bytes = (U + n_sh) * 3 * I * H * wb + (T / world) * H * 2 + T * H * 2
    else:
        kern = ("umma_fp8_prefill_kernel<gate_up+silu_mul+requant>" if fp8 and T * k / shape.E >= 48 else
                "umma_fp8_kernel<gate_up+silu_mul>" if fp8 else "umma_grouped_kernel<gate_up+silu_mul>")
        gu_bytes = (U + n_sh) * 2 * I * H * wb + rows * H * wb + rows * I * wb
    achieved = gu_bytes / gu_ms / 1e6
    gu_flops = rows * 2 * H * 2 * I
    # DRAM traffic of the same launch from the committed `ncu --set full` capture (profiles/ncu_traffic.json)
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        base = kern.split('<')[0]
        traffic = (tj.get(f"{base}:T={T}:{args.weights}") or tj.get(f"{base}<gate_up>:T={T}:{args.weights}")) \
            if world == 1 else None
    except (OSError, ValueError):
        pass
    layer_bytes = shape.algorithmic_bytes(T, U, weight_bytes=wb, fp8_scales=fp8) if world == 1 else None
    roofline = {"bound": "hbm", "kernel": kern, "achieved": achieved,
                "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                "peak_source": peak_src, "kernel_ms": gu_ms, "algorithmic_bytes_per_launch": gu_bytes,
                "peak_note": "hbm_gbs of MEASURED_PEAKS.json is a read+write copy; read-only bulk-copy streaming on the "
                             "same box type reaches 7.2 TB/s (profiles/r1d_ubench.jsonl), so frac can exceed 1",
                "stage_ms": stage_acc}
    if T * k / shape.E >= 96 and not fused:    # rows per local expert do not depend on the world size
        # prefill: the grouped GEMMs are a dense contraction bound by the tensor pipe
        roofline.update({"bound": "tensor", "achieved": gu_flops / gu_ms / 1e9, "peak": tc_peak, "unit": "TFLOP/s",
                         "frac": gu_flops / gu_ms / 1e9 / tc_peak, "algorithmic_flops_per_launch": gu_flops,
                         "peak_source": peak_src + (" x 2 for e4m3 (no measured FP8 peak)" if fp8 else ""),
                         "hbm": {"achieved": achieved, "peak": hbm_peak, "frac": achieved / hbm_peak}})
        fl = shape.flops_per_token() * T / world
        roofline["layer_tensor"] = {"algorithmic_flops": fl, "achieved": fl / ms_step / 1e9,
                                    "frac": fl / ms_step / 1e9 / tc_peak}
    if layer_bytes:
        roofline["layer"] = {"algorithmic_bytes": layer_bytes, "achieved": layer_bytes / ms_step / 1e6,
                             "frac": layer_bytes / ms_step / 1e6 / hbm_peak}
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        n = args.cpu_steps or (40 if T <= 64 else 8 if T <= 512 else 2)
        host_w = {k_: v.cpu() for k_, v in weights.items()}
        tps, ms, cores = _cpu_reference_run(shape, dt, host_w, T, n, 2)
        cpu_baseline = {"value": tps, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_step": ms,
                        "sample": f"{n} steps of the same workload (T={T}) after 2 warm-up, oracle port of the "
                                  "reference torch path (moe.py:128-176), torch threads = all host cores"}
    line = {"metric": METRIC, "value": T * 1e3 / ms_step, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if not fp8 else "fp8_e4m3 (fp32 block scales, fp32 accumulate)",
            "data": "synthetic",
            "config": {"workload": workload, "tokens_per_step": T, "experts_hit_mean": U,
                       "parallelism": (f"ep{world}-{args.ep_layout}" + (f"-sliced{args.ep_sliced}" if args.ep_sliced >= 0 else ""))
                       if world > 1 else "single",
                       "l2": f"{n_batches} distinct token batches rotated; weights touched per step "
                             f"({(U + 1) * 3 * H * I * 2 / 1e9:.2f} GB) exceed the 126 MB L2",
                       "expert_gemm_impl": impl_name},
            "e2e": e2e, "gpu_launches": launches, "clocks": clk.summary(), "roofline": roofline,
            "cpu_baseline": cpu_baseline}
    if sweep:
        line["sweep"] = sweep
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
