#!/usr/bin/env python
"""bench.py — GLM-4.5-Air MoE-layer tokens/s on B200 (BASELINE.json metric, configs[1]).

A "step" is one forward of the sparse-expert block over one batch of T synthetic tokens
(default T=32, the top of north_star's decode 1-32 range).  One JSON line on stdout:

  value      tokens/s with the token batches already resident in HBM (CUDA events over exactly K steps,
             max over ranks); `ms_per_step_median` = median of per-step event pairs taken in a second loop
  e2e        same metric through the module's host-buffer entry (pinned host in, pinned host
             out, H2D + D2H inside the timed region)
  roofline   HBM (decode) or tensor (prefill) roofline of the dominant kernel, timed live with CUDA
             events between stages (glm4moe_layer_forward_timed)
  cpu_baseline  the reference's own torch path on the box's host cores (N=1, rank 0)

`--impl reference` times the reference's CPU torch path on the same workload: the REAL reference
module from `oracle/_ref/` (installed by `oracle/make_ref.py`; `kind: "reference"`), or the oracle
port of it when that copy is absent (`kind: "port"`).  N>1: expert-parallel (experts sharded over
ranks, tokens replicated; the weighted expert rows are combined and reduced across ranks by one kernel
reading NVLink peer memory, `--ep-layout allreduce` = NCCL instead), strong scaling; one batch is
checked bit for bit against a single-GPU forward outside the timed region (`ep_bit_identical`).
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
READ_STREAM_GBS = 7200.0    # read-only bulk-copy streaming measured on this box type (profiles/r1d_ubench.jsonl)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], "measured (MEASURED_PEAKS.json)", d.get("bf16_tflops", 1650.0)
    return 6650.0, "fallback (B200_PROFILING.md)", 1590.0


def _fp8_peak(bf16_peak):
    """Dense e4m3 peak: measured by scripts/fp8_peak.py on the box (profiles/fp8_peak_r2.json) when
    committed, else 2 x the measured bf16 rate (stated as an assumption)."""
    p = os.path.join(ROOT, "profiles", "fp8_peak_r2.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["fp8_tflops"]), "measured (profiles/fp8_peak_r2.json)"
        except (OSError, ValueError, KeyError):
            pass
    return 2.0 * bf16_peak, "2 x the measured bf16 peak (no measured FP8 figure)"


class ClockSampler:
    """Samples SM clocks / throttle reasons during the timed region (NVML, ~1 ms period)."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.watts, self.limit_w = index, [], set(), [], None
        self.max_mhz, self._stop, self._thr = None, threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self.limit_w = pynvml.nvmlDeviceGetEnforcedPowerLimit(self._h) / 1000.0
        except Exception:
            self._nv = None

    NAMES = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
             "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}

    def sample_now(self):
        """One synchronous sample.  The main thread calls it right after the timed steps are ENQUEUED (the GPU is
        then still inside the timed region): a 3 ms region (20 steps at 8 GPUs) otherwise races the sampler thread,
        which has to win the GIL from the launching thread."""
        nv = self._nv
        if nv is None:
            return
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
            self.watts.append(nv.nvmlDeviceGetPowerUsage(self._h) / 1000.0)   # NVML averages this over ~0.1 s
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
            for n, bit in self.NAMES.items():
                if r & bit:
                    self.reasons.add(n)
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self.sample_now()
            time.sleep(0.001)

    def start(self):
        if self._nv is not None and self._thr is None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
            time.sleep(0.005)      # the thread is up and sampling before the caller goes on
        return self

    def mark(self):
        """Samples taken so far are outside the timed region: forget them."""
        self.samples.clear()
        self.reasons.clear()
        self.watts.clear()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples),
                "power_w_max": round(max(self.watts), 1) if self.watts else None, "power_limit_w": self.limit_w}


def _hf_config(shape):
    from transformers.models.glm4_moe import Glm4MoeConfig
    return Glm4MoeConfig(hidden_size=shape.H, n_routed_experts=shape.E, num_experts_per_tok=shape.k,
                         moe_intermediate_size=shape.I, n_shared_experts=1, n_group=1, topk_group=1,
                         routed_scaling_factor=shape.routed_scaling_factor, norm_topk_prob=True)


def _workload(shape, T, fp8):
    regime = "decode batch" if T <= 256 else "prefill"
    return (f"{shape.name} MoE layer (H={shape.H}, E={shape.E}, top-{shape.k}, I={shape.I}, 1 shared) "
            f"{'FP8 e4m3 block-128 expert weights (power-of-two scales), bf16 activations in/out,' if fp8 == 'fp8_pow2' else 'FP8 e4m3 block-128 expert weights, bf16 activations in/out,' if fp8 else 'bf16'} "
            f"random-init, {regime} T={T}")


def _dominant_kernel(shape, T, U, world, fused, fp8, pow2):
    """(name, algorithmic bytes of one launch, its flops).  Dominant kernel = the gate_up grouped GEMM (decode:
    the fused gate_up + down + combine launch).  Bytes = weights of the (U local + 1 shared) experts it touches
    + the activation rows it reads + the rows it writes."""
    H, I, k = shape.H, shape.I, shape.k
    wb = 1 if fp8 else 2          # bytes per expert-weight element (and per gathered / activation row element)
    rows = T * k / world + T / world
    if fused:
        # decode regime: gate_up + SiLU*mul + down + combine are ONE persistent launch (decode_fused.cu)
        kern = "decode_fused_kernel<gate_up+silu_mul+down+combine>"
        by = (U + 1) * 3 * I * H * wb + (T / world) * H * 2 + T * H * 2
    else:
        kern = ("umma_fp8_mx_kernel<gate_up+silu_mul+requant> (tcgen05 block-scaled MMA)" if pow2 and T * k / shape.E >= 48 else
                "umma_fp8_prefill_kernel<gate_up+silu_mul+requant>" if fp8 and T * k / shape.E >= 48 else
                "umma_fp8_kernel<gate_up+silu_mul>" if fp8 else "umma_grouped_kernel<gate_up+silu_mul>")
        by = (U + 1) * 2 * I * H * wb + rows * H * wb + rows * I * wb
    return kern, by, rows * 2 * H * 2 * I


def _config(shape, T, fp8, n_batches):
    """The `config` object: identical in both arms (the driver compares them)."""
    return {"workload": _workload(shape, T, fp8), "tokens_per_step": T,
            "l2": f"{n_batches} distinct token batches rotated; expert weights touched per step exceed the 126 MB L2 "
                  "from T >= 2 (no L2 flush between steps)"}


class _CpuReference:
    """The reference's CPU torch path on the host cores: the real module when oracle/_ref exists."""

    def __init__(self, shape, dt, weights):
        from oracle import reference_live as R
        self.shape, self.weights = shape, weights
        self.kind = "reference" if R.available() else "port"
        self.what = ("the reference's own Glm4MoeMoE.forward (moe.py:128-176) imported from oracle/_ref "
                     "(debug file read patched out)") if self.kind == "reference" else \
            "oracle port of the reference torch path (moe.py:128-176); oracle/_ref not installed"
        self.module = None
        if self.kind == "reference":
            self.module = R.build_reference_module(shape, dt)
            R.fill_reference_module(self.module, weights)

    def __call__(self, x):
        if self.module is not None:
            return self.module(x)
        from oracle import moe_oracle as O
        return O.moe_forward(x, self.weights, self.shape.k, self.shape.routed_scaling_factor)["out"]


def _cpu_reference_run(shape, dt, weights, T, steps, warmup, budget_s=150.0):
    """Returns (tok/s, ms/step, cores, steps actually timed, kind, what)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = _CpuReference(shape, dt, weights)
    xs = [synth.make_tokens(shape, T, dt, b) for b in range(min(N_BATCHES, steps + warmup))]
    with torch.inference_mode():
        for i in range(warmup):
            ref(xs[i % len(xs)])
        t0 = time.perf_counter()
        done = 0
        for i in range(steps):
            ref(xs[(warmup + i) % len(xs)])
            done += 1
            if time.perf_counter() - t0 > budget_s:      # bounded sample: a prefill step takes seconds on a CPU
                break
        dt_s = time.perf_counter() - t0
    return T * done / dt_s, dt_s / done * 1e3, cores, done, ref.kind, ref.what


def _timed(fn, xs, warm, n):
    """ms per call of fn over the rotating device batches xs (CUDA events on the current stream)."""
    for i in range(warm):
        fn(xs[i % len(xs)])
    torch.cuda.synchronize()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for i in range(n):
        fn(xs[i % len(xs)])
    s1.record()
    torch.cuda.synchronize()
    return s0.elapsed_time(s1) / n


def _context_legs(moe, shape, dt, dev, weights, T, ms_step, hbm_peak, tc_peak):
    """Short extra measurements printed beside the headline (N = 1, bf16 arm only); none of them is `value`.
      gpu_reference  the reference's own Glm4MoeMoE.forward (moe.py:128-176: per-expert cuBLAS F.linear loop, imported
                     unmodified from oracle/_ref) on THIS GPU with the same weights and batches: SURVEY 2.1's
                     "kernel to beat" on Blackwell, where the headline ratio is against the 16-core CPU run
      other_configs  the other BASELINE.json configurations on the same layer: T = 1 decode, T = 4096 prefill (bf16,
                     same module) and the e4m3 tier at T = 32 (configs[2]; a second module quantised from the same
                     weights).  Same timing rules: device batches rotated, CUDA events, warm-up first
    Every leg is fenced: a failure becomes {"error": ...} and never costs the headline line."""
    out = {"gpu_reference": None, "other_configs": []}
    with torch.inference_mode():
        try:
            from oracle import reference_live as R
            if R.available():
                ref = R.build_reference_module(shape, dt, device=str(dev))
                R.fill_reference_module(ref, weights)
                xs = [synth.make_tokens(shape, T, dt, b).to(dev) for b in range(8)]
                ms = _timed(ref, xs, 3, 20)
                got, want = moe(xs[0]).float(), ref(xs[0]).float()
                rel = float((got - want).norm() / want.norm())
                out["gpu_reference"] = {
                    "what": "reference Glm4MoeMoE.forward (oracle/_ref, unmodified) on the same GPU, same weights, "
                            f"T={T}, 20 steps after 3 warm-up, CUDA events",
                    "ms_per_step": ms, "tokens_per_s": T / ms * 1e3, "speedup_device_timed": ms / ms_step,
                    "rel_l2_ours_vs_reference": rel}
                del ref
                torch.cuda.empty_cache()
            else:
                out["gpu_reference"] = {"unavailable": "oracle/_ref not installed (python oracle/make_ref.py)"}
        except Exception as e:  # noqa: BLE001
            out["gpu_reference"] = {"error": f"{type(e).__name__}: {e}"[:300]}

        def leg(module, Ts, wfmt, peak_tc):
            xs = [synth.make_tokens(shape, Ts, dt, 100 + b).to(dev) for b in range(8 if Ts <= 256 else 4)]
            ms = _timed(module, xs, 5, 100 if Ts <= 256 else 20)
            us = []
            for x in xs:
                module(x)
                us.append(int(torch.unique(module.workspace_view("idx", Ts)).numel()))
            U = sum(us) / len(us)
            fp8_ = wfmt != "bf16"
            by = shape.algorithmic_bytes(Ts, U, weight_bytes=1 if fp8_ else 2, fp8_scales=fp8_)
            fl = shape.flops_per_token() * Ts
            tensor_bound = Ts * shape.k / shape.E >= 96
            return {"workload": f"{shape.name} MoE layer, {wfmt}, T={Ts}", "tokens": Ts, "weights": wfmt,
                    "ms_per_step": ms, "tokens_per_s": Ts / ms * 1e3, "experts_hit_mean": U,
                    "layer_GBps": by / ms / 1e6, "layer_TFLOPs": fl / ms / 1e9,
                    "bound": "tensor" if tensor_bound else "hbm",
                    "layer_frac": (fl / ms / 1e9 / peak_tc) if tensor_bound else (by / ms / 1e6 / hbm_peak),
                    "peak": peak_tc if tensor_bound else hbm_peak, "impl": module.last_impl,
                    "fused_decode_kernel": module.last_fused}

        for Ts in (1, 4096):
            try:
                out["other_configs"].append(leg(moe, Ts, "bf16", tc_peak))
            except Exception as e:  # noqa: BLE001
                out["other_configs"].append({"workload": f"bf16 T={Ts}", "error": f"{type(e).__name__}: {e}"[:300]})
        try:
            torch.set_default_dtype(dt)
            torch.set_default_device(dev)
            try:
                q = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=256)
            finally:
                torch.set_default_device("cpu")
                torch.set_default_dtype(torch.float32)
            q.load_local_stacks(weights)
            q.quantize_fp8(drop_native=True, pow2_scales=False)
            out["other_configs"].append(leg(q, 32, "fp8_e4m3 (fp32 block scales)", _fp8_peak(tc_peak)[0]))
            del q
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001
            out["other_configs"].append({"workload": "fp8 T=32", "error": f"{type(e).__name__}: {e}"[:300]})
    return out


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
    ap.add_argument("--no-context", action="store_true",
                    help="skip the context legs (reference module on the GPU, the other BASELINE configurations)")
    ap.add_argument("--ep-layout", default="p2p", choices=["p2p", "allreduce", "alltoall"],
                    help="expert-parallel layout for --gpus > 1 (distributed.py): p2p = combine + reduction in "
                         "one kernel over NVLink peer memory, allreduce / alltoall = NCCL")
    ap.add_argument("--weights", default="bf16", choices=["bf16", "fp8", "fp8_pow2"],
                    help="fp8 = e4m3 expert weights with 128x128 fp32 block scales, per-token 1x128 activations "
                         "(BASELINE configs[2]); single GPU")
    ap.add_argument("--ep-sliced", type=int, default=-1, choices=[-1, 0, 1],
                    help="p2p layout: -1 pick by size, 0 every rank pulls every row, 1 reduce the own token slice "
                         "and store it to the peers (glm4moe_set_ep_sliced)")
    ap.add_argument("--no-ep-check", action="store_true", help="skip the single-GPU bit-identity check at N > 1")
    ap.add_argument("--one-forward", action="store_true",
                    help="after warm-up run ONE forward of batch 0 and exit (the launch an ncu capture profiles)")
    ap.add_argument("--knob", action="append", default=[], metavar="K=V",
                    help="glm4moe_debug_set(K, V) before the run (A/B measurements; include/glm4_moe_b200.h)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    shape = synth.SHAPES[args.shape]
    dt = torch.bfloat16
    T = args.tokens
    fp8 = args.weights in ("fp8", "fp8_pow2")
    pow2 = args.weights == "fp8_pow2"      # UE8M0 scales: prefill GEMMs with hardware block scaling (umma_fp8_mx.cu)
    if fp8 and world > 1:
        raise SystemExit("--weights fp8 is a single-GPU bench")
    n_batches = N_BATCHES if T <= 1024 else 4
    config = _config(shape, T, args.weights if fp8 else False, n_batches)
    if args.knob:
        config["knobs"] = list(args.knob)

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return
        weights = synth.make_weights(shape, dt)
        steps, warm = max(1, args.steps), max(1, args.warmup)
        tps, ms, cores, done, kind, what = _cpu_reference_run(shape, dt, weights, T, steps, warm)
        line = {"metric": METRIC, "value": tps, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
                "steps": done, "warmup": warm, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": config,
                "cpu_baseline": {"value": tps, "unit": UNIT, "cores": cores, "kind": kind,
                                 "sample": f"{done} steps of the full workload after {warm} warm-up: {what}; "
                                           f"torch threads = all {cores} host cores"},
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

    for kv in args.knob:
        from glm4_moe_b200 import _lib
        k_, v_ = kv.split("=")
        _lib.check(_lib.load().glm4moe_debug_set(int(k_), int(v_)))

    torch.set_default_dtype(dt)
    torch.set_default_device(dev)
    moe = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=max(T, 256), ep_rank=rank, ep_size=world,
                                   peer_workspace=(world > 1 and args.ep_layout == "p2p"))
    torch.set_default_device("cpu")
    torch.set_default_dtype(torch.float32)
    weights = synth.make_weights_device(shape, dt, dev, experts=range(moe.expert_begin, moe.expert_end))
    moe.load_local_stacks(weights)
    if fp8:
        moe.quantize_fp8(drop_native=True, pow2_scales=pow2)      # offline conversion (fp8.quantize_block128), outside the timed region
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

    clk = ClockSampler(local_rank)
    with torch.inference_mode():
        for i in range(max(args.warmup, 3)):
            step(batches[i % n_batches])
        torch.cuda.synchronize()
        if args.one_forward:
            # ncu target (scripts/ncu_traffic.sh): the forward after the warm-up is the captured one
            step(batches[0])
            torch.cuda.synchronize()
            idx0 = moe.workspace_view("idx", T)
            U0 = int(torch.unique(idx0[(idx0 >= moe.expert_begin) & (idx0 < moe.expert_end)]).numel())
            kern, by, fl = _dominant_kernel(shape, T, U0, world, moe.last_fused, fp8, pow2)
            print(json.dumps({"one_forward": f"{shape.name}:T={T}:{args.weights}:n={world}", "kernel": kern,
                              "experts_hit": U0, "algorithmic_bytes": by, "flops": fl,
                              "warmup_forwards": max(args.warmup, 3)}))
            return
        # ---- timed region: exactly K steps, CUDA events, max over ranks.  Everything host-side that
        # differs per rank (NVML init, thread start) happens BEFORE the barrier; a tiny NCCL all-reduce
        # enqueued right in front of the start event is the cross-rank *device* barrier, so no rank's clock
        # starts while a peer is still behind (a p2p step spins on its peers' flags). ----
        clk.start()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_t = torch.zeros(1, device=dev)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            dist.all_reduce(sync_t)
        clk.mark()
        start.record()
        for i in range(args.steps):
            step(batches[i % n_batches])
        end.record()
        clk.sample_now()              # the queue is still draining: inside the timed region
        torch.cuda.synchronize()
        clocks = clk.summary()
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

        # per-step event pairs (second loop: an event between steps costs ~2-3 us and breaks the
        # programmatic overlap of consecutive forwards, so it stays out of the headline region)
        n_med = min(args.steps, 100)
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_med + 1)]
        if world > 1:
            dist.barrier()
            dist.all_reduce(sync_t)
        evs[0].record()
        for i in range(n_med):
            step(batches[i % n_batches])
            evs[i + 1].record()
        torch.cuda.synchronize()
        per_step = [evs[i].elapsed_time(evs[i + 1]) for i in range(n_med)]
        ms_median = statistics.median(per_step)
        if world > 1:
            t = torch.tensor([ms_median], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_median = float(t.item())

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
        # expert-parallel load: local experts hit per rank per batch (the step waits for the slowest rank;
        # SURVEY 8e "per-rank load imbalance (max / mean)")
        ep_load = None
        if world > 1:
            mine = torch.tensor(hits, dtype=torch.float32, device=dev)
            allh = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allh, mine)
            allh = torch.stack(allh)                                   # [rank, batch]
            ep_load = {"experts_hit_per_rank_mean": [round(v, 2) for v in allh.mean(1).tolist()],
                       "max_over_mean_per_step": round(float((allh.max(0)[0] / allh.mean(0)).mean()), 3)}

        # N > 1: one batch of the expert-parallel layer against a single-GPU forward, bit for bit
        ep_bit_identical = None
        if world > 1 and not args.no_ep_check:
            y_ep = layer(batches[0]).clone()
            ok = torch.ones(1, device=dev)
            if rank == 0:
                torch.set_default_dtype(dt)
                torch.set_default_device(dev)
                single = glm4_moe_b200.Glm4MoeMoE(_hf_config(shape), max_tokens=max(T, 256))
                torch.set_default_device("cpu")
                torch.set_default_dtype(torch.float32)
                single.load_local_stacks(synth.make_weights_device(shape, dt, dev))
                y_1 = single(batches[0])
                ok[0] = 1.0 if torch.equal(y_1, y_ep) else 0.0
                ok_close = bool(((y_1.float() - y_ep.float()).abs().max() <=
                                 2e-2 * y_1.float().abs().max()).item())
                del single
            dist.broadcast(ok, 0)
            ep_bit_identical = bool(ok.item() == 1.0)
            if rank == 0 and not ep_bit_identical and args.ep_layout == "p2p":
                print(f"WARNING: expert-parallel output differs from the single-GPU forward (close: {ok_close})",
                      file=sys.stderr)

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
                by = shape.algorithmic_bytes(Ts, Us, weight_bytes=1 if fp8 else 2, fp8_scales=fp8)
                fl = shape.flops_per_token() * Ts
                sweep.append({"T": Ts, "ms": ms, "tokens_per_s": Ts / ms * 1e3, "experts_hit": Us,
                              "GBps": by / ms / 1e6, "TFLOPs": fl / ms / 1e9,
                              "impl": moe.last_impl, "fused_decode_kernel": moe.last_fused,
                              "tile_rows": _lib_tile_rows(moe, Ts)})
    clk.stop()

    if world > 1 and args.ep_layout == "p2p":
        layer.close_p2p()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, peak_src, tc_peak = _peaks()
    tc_src = peak_src
    if fp8:
        tc_peak, tc_src = _fp8_peak(tc_peak)
    H, I, k = shape.H, shape.I, shape.k
    wb = 1 if fp8 else 2
    gu_ms = stage_acc["gate_up"]     # decode: this stage interval brackets exactly the fused launch
    kern, gu_bytes, gu_flops = _dominant_kernel(shape, T, U, world, fused, fp8, pow2)
    achieved = gu_bytes / gu_ms / 1e6
    # DRAM traffic of this launch as ncu saw it on THIS bench's batch 0 (`bench.py --one-forward` under
    # `ncu --set full`, summarised by scripts/ncu_traffic.py into profiles/ncu_traffic_r2.json together
    # with the experts that launch hit); null when no capture of this exact workload is committed.
    traffic, traffic_info = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic_r2.json")))
        ent = tj.get(f"{shape.name}:T={T}:{args.weights}:n={world}")
        if ent and ent.get("kernel", "").split("<")[0] == kern.split("<")[0]:
            traffic = ent["dram_bytes"]
            traffic_info = {k_: ent[k_] for k_ in ("experts_hit", "algorithmic_bytes", "duration_us", "capture")
                            if k_ in ent}
    except (OSError, ValueError):
        pass
    layer_bytes = shape.algorithmic_bytes(T, U, weight_bytes=wb, fp8_scales=fp8) if world == 1 else None
    roofline = {"bound": "hbm", "kernel": kern, "achieved": achieved,
                "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                "traffic_capture": traffic_info,
                "peak_source": peak_src, "kernel_ms": gu_ms, "algorithmic_bytes_per_launch": gu_bytes,
                "frac_of_read_stream": achieved / READ_STREAM_GBS, "read_stream_peak": READ_STREAM_GBS,
                "peak_note": "`peak` (hbm_gbs of MEASURED_PEAKS.json) is a read+write copy; this kernel only reads, and "
                             "read-only bulk-copy streaming on the same box type reaches 7.2 TB/s "
                             "(profiles/r1d_ubench.jsonl): `frac_of_read_stream` is the fraction of that ceiling",
                "stage_ms": stage_acc}
    if T * k / shape.E >= 96 and not fused:    # rows per local expert do not depend on the world size
        # prefill: the grouped GEMMs are a dense contraction bound by the tensor pipe
        roofline.update({"bound": "tensor", "achieved": gu_flops / gu_ms / 1e9, "peak": tc_peak, "unit": "TFLOP/s",
                         "frac": gu_flops / gu_ms / 1e9 / tc_peak, "algorithmic_flops_per_launch": gu_flops,
                         "peak_source": tc_src,
                         "hbm": {"achieved": achieved, "peak": hbm_peak, "frac": achieved / hbm_peak}})
        fl = shape.flops_per_token() * T / world
        roofline["layer_tensor"] = {"algorithmic_flops": fl, "achieved": fl / ms_step / 1e9,
                                    "frac": fl / ms_step / 1e9 / tc_peak}
    if layer_bytes:
        roofline["layer"] = {"algorithmic_bytes": layer_bytes, "achieved": layer_bytes / ms_step / 1e6,
                             "frac": layer_bytes / ms_step / 1e6 / hbm_peak,
                             "frac_of_read_stream": layer_bytes / ms_step / 1e6 / READ_STREAM_GBS}
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        n = args.cpu_steps or (40 if T <= 64 else 8 if T <= 512 else 2)
        host_w = {k_: v.cpu() for k_, v in weights.items()}
        tps, ms, cores, done, kind, what = _cpu_reference_run(shape, dt, host_w, T, n, 2, budget_s=60.0)
        cpu_baseline = {"value": tps, "unit": UNIT, "cores": cores, "kind": kind, "ms_per_step": ms,
                        "sample": f"{done} steps of the same workload (T={T}) after 2 warm-up: {what}; "
                                  f"torch threads = all {cores} host cores"}
    context = None
    if world == 1 and not fp8 and not args.no_context and T <= 256:
        context = _context_legs(moe, shape, dt, dev, weights, T, ms_step, hbm_peak, tc_peak)
    line = {"metric": METRIC, "value": T * 1e3 / ms_step, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "ms_per_step_median": ms_median,
            "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if not fp8 else
            "fp8_e4m3 (power-of-two block scales applied by the MMA, fp32 accumulate)" if pow2 else
            "fp8_e4m3 (fp32 block scales, fp32 accumulate)",
            "data": "synthetic", "config": config,
            "details": {"experts_hit_mean": U, "expert_gemm_impl": impl_name, "fused_decode_kernel": bool(fused),
                        "parallelism": (f"ep{world}-{args.ep_layout}" + (f"-sliced{args.ep_sliced}" if args.ep_sliced >= 0 else ""))
                        if world > 1 else "single",
                        "weights_touched_per_step_gb": (U + 1) * 3 * H * I * wb / 1e9,
                        "ep_bit_identical": ep_bit_identical, "ep_load": ep_load},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu_baseline}
    if ep_bit_identical is not None:
        line["ep_bit_identical"] = ep_bit_identical
    if sweep:
        line["sweep"] = sweep
    if context:
        line.update(context)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _lib_tile_rows(moe, T):
    from glm4_moe_b200 import _lib
    return int(_lib.load().glm4moe_layer_tile_rows(moe._handle, T)) if moe._handle else None


if __name__ == "__main__":
    main()
