#!/usr/bin/env python
"""Benchmark of the one hot path: the worker's (sharded) decoder forward.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  A "step" is one decode step: one new token for
every sequence of the batch, with the KV cache already holding `ctx` tokens.

N = 1 : config[2] of BASELINE.json — Llama-3-8B shape, bf16, random-init weights,
        batch 1, 2k-token context, paged KV (the north-star target).  Also reports
        the 2k-token prefill and (extra) batch-32 decode.
N > 1 : the same model split into N contiguous layer ranges (reference sharder
        formula), one stage per GPU, hidden states handed over with NCCL
        send/recv, N micro-batches of `batch` sequences in flight (weak scaling:
        per-GPU batch fixed, per-GPU weights shrink).

value  : decode tokens/s with everything resident in HBM (CUDA-graph replay).
e2e    : the same metric through the host-facing call: every step copies the
         previous tokens host->device from pinned memory, runs the step, copies the
         new tokens device->host and synchronises.
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
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "decode_tokens_per_s"
UNIT = "tokens/s"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Polls SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index=0, period_s=0.02):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self.period = period_s
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.nv, self.err = None, str(e)

    def _names(self, mask):
        nv, out = self.nv, []
        table = [("hw_slowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                 ("hw_thermal_slowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                 ("sw_thermal_slowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                 ("sw_power_cap", "nvmlClocksThrottleReasonSwPowerCap"),
                 ("hw_power_brake", "nvmlClocksThrottleReasonHwPowerBrakeSlowdown")]
        for name, attr in table:
            bit = getattr(nv, attr, None)
            if bit is not None and mask & bit:
                out.append(name)
        return out

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                self.reasons.update(self._names(fn(self.h)))
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "err", "nvml unavailable")}
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def decode_step_bytes(spec, n_layers, has_head, B, S):
    """Algorithmic HBM bytes of one decode stage pass."""
    kv = spec.kv_bytes_per_token(n_layers)
    return n_layers * spec.layer_weight_bytes() + (spec.head_weight_bytes() if has_head else 0) + B * S * kv + B * kv


def prefill_flops(spec, B, T):
    """Algorithmic FLOPs of a prefill: linear + causal attention + last-position head."""
    p_layer = spec.layer_weight_bytes() // 2
    return 2 * spec.n_layers * p_layer * B * T + 2 * spec.d_model * spec.vocab * B \
        + 4 * B * (T * T // 2) * spec.n_heads * spec.head_dim * spec.n_layers


def ev_time_ms(fn, stream=None):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


# ---------------------------------------------------------------------------------- ours, 1 GPU
def run_single(args, spec, peaks):
    from dli_b200 import ops
    from dli_b200.engine import ShardEngine

    B, ctx, K, W = args.batch, args.ctx, args.steps, args.warmup
    max_len = ctx + 2 * (K + W) + 16
    eng = ShardEngine(spec, max_batch=B, max_seq_len=max_len, page_size=args.page_size)
    eng.load_random(seed=0)
    g = torch.Generator().manual_seed(1)
    prompt = torch.randint(0, spec.vocab, (B, ctx), generator=g, dtype=torch.int32).pin_memory()

    # ---- prefill (also fills the KV cache with `ctx` tokens per sequence) -----------------
    def do_prefill():
        eng.start_batch(B, ctx, max_len)
        ids = prompt.to("cuda", non_blocking=True).reshape(-1)
        eng.prefill(B, ctx, ids=ids)

    do_prefill()  # warm-up (function attributes, workspace)
    pf_ms = min(ev_time_ms(do_prefill) for _ in range(2))
    pf_flops = prefill_flops(spec, B, ctx)

    # ---- decode, device-resident -------------------------------------------------------------
    for _ in range(max(W, 3)):
        eng.decode_step(B)
    torch.cuda.synchronize()
    n0 = ops.launch_count
    with ClockSampler() as clk:
        ms = ev_time_ms(lambda: [eng.decode_step(B) for _ in range(K)])
    launches = ops.launch_count - n0
    ms_step = ms / K
    S_mid = ctx + max(W, 3) + K // 2
    step_bytes = decode_step_bytes(spec, spec.n_layers, True, B, S_mid)
    value = B * K / (ms / 1e3)

    # ---- decode through the host-facing call -----------------------------------------------
    t_in = torch.zeros(B, dtype=torch.int32).pin_memory()
    t_out = torch.zeros(B, dtype=torch.int32).pin_memory()
    t_in.copy_(eng._state(B)["tokens"].cpu())
    for _ in range(3):
        eng.decode_step_host(t_in, t_out)
        t_in.copy_(t_out)

    def host_loop():
        for _ in range(K):
            eng.decode_step_host(t_in, t_out)
            t_in.copy_(t_out)

    e2e_ms = ev_time_ms(host_loop)
    e2e = {"value": B * K / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": 4 * B, "d2h_bytes_per_step": 4 * B,
           "ms_per_step": e2e_ms / K}

    # ---- dominant kernel: the gate/up weight-streaming projection ----------------------------
    roof = dominant_kernel_roofline(eng, spec, B, peaks)
    roof_step = {"bound": "hbm", "achieved": step_bytes / (ms_step / 1e3) / 1e9, "peak": peaks["hbm_gbs"],
                 "unit": "GB/s", "frac": step_bytes / (ms_step / 1e3) / 1e9 / peaks["hbm_gbs"],
                 "bytes_per_step": step_bytes, "frac_of_nominal_8TBps": step_bytes / (ms_step / 1e3) / 8e12}

    extra = {
        "prefill_tokens_per_s": B * ctx / (pf_ms / 1e3), "prefill_ms": pf_ms,
        "prefill_roofline": {"bound": "tensor", "achieved": pf_flops / (pf_ms / 1e3) / 1e12,
                             "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
                             "frac": pf_flops / (pf_ms / 1e3) / 1e12 / peaks["bf16_tflops"]},
        "roofline_step": roof_step, "attn_splits": eng._state(B)["splits"],
        "weights_gb": eng.weight_bytes() / 1e9, "peaks": peaks["source"],
    }
    if args.batch == 1 and not args.no_extra:
        try:
            extra["decode_b32"] = batch_decode_extra(spec, 32, ctx, min(K, 32), peaks, args.page_size, eng)
        except Exception as e:  # noqa: BLE001
            extra["decode_b32"] = {"error": str(e)[:200]}
    return {"value": value, "ms_per_step": ms_step, "e2e": e2e, "roofline": roof, "gpu_launches": launches,
            "clocks": clk.summary(), "extra": extra,
            "config": {"workload": f"{args.model} bf16 random-init, 1 GPU, batch {B} greedy decode at {ctx}-token context, "
                                   f"paged KV (page {args.page_size}); BASELINE.json configs[2]",
                       "batch": B, "ctx": ctx, "page_size": args.page_size, "parallelism": "1 stage",
                       "l2": f"inputs larger than L2: {step_bytes / 1e9:.2f} GB streamed per step vs 126 MB L2"}}


def dominant_kernel_roofline(eng, spec, B, peaks):
    """Time the dominant decode kernel alone with CUDA events on its stream: the gated
    gate/up projection of every layer back to back (each launch reads a different
    layer's 2*d*ff*2 bytes, so nothing is L2-resident)."""
    from dli_b200 import ops

    x = torch.randn(B, spec.d_model, device="cuda").to(torch.bfloat16)
    out = torch.empty(B, spec.d_ff, device="cuda", dtype=torch.bfloat16)
    ws = [L["wgu"] for L in eng.layers]

    def loop():
        for w in ws:
            ops.linear_gated_silu(x, w, out=out)

    loop()
    ms = min(ev_time_ms(loop) for _ in range(3)) / len(ws)
    nbytes = 2 * spec.d_ff * spec.d_model * 2
    ach = nbytes / (ms / 1e3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("gate_up_projection_dram_bytes_per_launch")
    return {"bound": "hbm", "kernel": "gemv_kernel<M,gated> (gate/up projection)" if B <= 4 else
            "gemm_tcgen05_kernel<swap,gated> (gate/up projection)", "achieved": ach, "peak": peaks["hbm_gbs"],
            "unit": "GB/s", "frac": ach / peaks["hbm_gbs"], "traffic": traffic, "bytes_per_launch": nbytes,
            "us_per_launch": ms * 1e3, "peak_source": peaks["source"]}


def batch_decode_extra(spec, B, ctx, K, peaks, page_size, old_eng):
    """Extra (not the headline): batch-32 decode at the same context."""
    from dli_b200.engine import ShardEngine

    max_len = ctx + K + 8
    eng = ShardEngine(spec, max_batch=B, max_seq_len=max_len, page_size=page_size)
    eng.w, eng.layers = old_eng.w, old_eng.layers  # share the weights already in HBM
    g = torch.Generator().manual_seed(2)
    ids = torch.randint(0, spec.vocab, (B * ctx,), generator=g, dtype=torch.int32).cuda()
    eng.start_batch(B, ctx, max_len)
    pf_ms = ev_time_ms(lambda: eng.prefill(B, ctx, ids=ids))
    for _ in range(3):
        eng.decode_step(B)
    ms = ev_time_ms(lambda: [eng.decode_step(B) for _ in range(K)]) / K
    nbytes = decode_step_bytes(spec, spec.n_layers, True, B, ctx + 3 + K // 2)
    out = {"tokens_per_s": B / (ms / 1e3), "ms_per_step": ms, "hbm_frac": nbytes / (ms / 1e3) / 1e9 / peaks["hbm_gbs"],
           "bytes_per_step": nbytes, "prefill_tokens_per_s": B * ctx / (pf_ms / 1e3),
           "prefill_tensor_frac": prefill_flops(spec, B, ctx) / (pf_ms / 1e3) / 1e12 / peaks["bf16_tflops_sustained"]}
    eng.w, eng.layers = {}, []
    eng.unload()
    return out


# ---------------------------------------------------------------------------------- CPU baseline
def _use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 for multi-process launches; the CPU arm is meant to use
    every host core this process may run on."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    if os.environ.get("OMP_NUM_THREADS") and torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_baseline(spec, args, new_tokens=4, prompt_len=16):
    from oracle import ref_worker_port as P

    _use_all_host_threads()
    t0 = time.time()
    r = P.time_cpu_decode(spec, prompt_len, new_tokens, greedy=True, warmup_tokens=0)
    return {"value": r["decode_tok_s"], "unit": UNIT, "cores": r["threads"], "kind": "port",
            "sample": r["sample"] + f" (transformers generate, the reference's worker/app.py path with USE_GPU=0); "
                                    f"host cpu_count={os.cpu_count()}", "wall_s": time.time() - t0,
            "prefill_tok_s": r["prefill_tok_s"]}


def run_reference(args, spec):
    """--impl reference: the reference's CPU path (port, see oracle/ref_worker_port.py),
    all host threads, each step = one decoded token at a short context (bounded sample)."""
    from oracle import ref_worker_port as P

    _use_all_host_threads()
    model = P.build_cpu_model(spec)
    tok = P.IdTokenizer()
    prompt_len = 16
    g = torch.Generator().manual_seed(1)
    prompt = " ".join(str(int(i)) for i in torch.randint(0, spec.vocab, (prompt_len,), generator=g))
    K, W = min(args.steps, 8), min(args.warmup, 1)
    if W:
        P.run_inference_port(model, tok, prompt, prompt_len + 1 + W, timeout=1e9, greedy=True)
    t0 = time.time()
    P.run_inference_port(model, tok, prompt, prompt_len + 1, timeout=1e9, greedy=True)
    t_pf = time.time() - t0
    t0 = time.time()
    P.run_inference_port(model, tok, prompt, prompt_len + 1 + K, timeout=1e9, greedy=True)
    dt = max(time.time() - t0 - t_pf, 1e-9)
    v = K / dt
    sample = (f"{args.model} bf16 on CPU via transformers generate (reference worker/app.py path, USE_GPU=0), greedy; "
              f"bounded sample: batch 1, {prompt_len}-token prompt, {K} decoded tokens (prefill time subtracted); "
              f"host cpu_count={os.cpu_count()}")
    return {"metric": METRIC, "value": v, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": K,
            "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{args.model} bf16 random-init, batch 1 greedy decode (CPU bounded sample)"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# ---------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama-3-8b")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--page-size", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    args = ap.parse_args()

    from dli_b200.engine import NAMED_SPECS

    spec = NAMED_SPECS[args.model]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args, spec)), flush=True)
        return

    peaks = measured_peaks()
    if world > 1 or args.gpus > 1:
        from dli_b200.pipeline import run_pipeline_bench

        res = run_pipeline_bench(args, spec, peaks)
        if rank != 0:
            return
    else:
        res = run_single(args, spec, peaks)
    line = {"metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": res["config"],
            "e2e": res["e2e"], "roofline": res["roofline"], "gpu_launches": res["gpu_launches"],
            "clocks": res["clocks"], "extra": res.get("extra", {})}
    if args.gpus == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_baseline(spec, args)
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": f"failed: {str(e)[:200]}"}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
