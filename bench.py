#!/usr/bin/env python
"""Benchmark of the one hot path: the worker's (sharded) decoder forward.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  A "step" is one decode step: one new token for
every sequence of the batch, with the KV cache already holding `ctx` tokens.

N = 1 : config[2] of BASELINE.json — Llama-3-8B shape, bf16, random-init weights,
        batch 1, 2k-token context, paged KV (the north-star target).  Also reports
        the 2k-token prefill and (extra) batch-32 decode.
N > 1 : the same model split into N contiguous layer ranges (reference sharder
        formula), one stage per GPU; decode hands hidden states, head candidates and
        tokens over in per-tick mailboxes (one NCCL all-gather per tick, issued on its
        own stream under the next tick's kernels), 2(N+1) micro-batch slots of `batch`
        sequences in flight so that every stage works every tick (weak scaling: per-GPU
        batch fixed, per-GPU weights shrink).  Before timing, rank 0 checks the
        pipeline's first tokens against the single-GPU engine on the same weights and
        the run aborts on a mismatch (extra.parity_check); extra.named_config runs the
        BASELINE.json configuration that is stated for exactly N GPUs.

value  : decode tokens/s with everything resident in HBM (CUDA-graph replay).
e2e    : the same metric through the plugin call the reference's master makes
         (reference worker/app.py:254-330 `run_inference`): POST /inference on the
         drop-in Worker (werkzeug test client, no socket) with the 2k-token prompt as
         host text -> tokenise -> host->device copy -> prefill -> K decode steps ->
         device->host copy of the tokens -> decode to text -> JSON.  Decode rate =
         K / (t[request with K+1 new tokens] - t[request with 1 new token]); the whole
         request's seconds are reported beside it.
extra.hf_gpu : transformers' own generate (bf16, sdpa) on the same GPU and shape — the only
         existing GPU implementation of this path (what the reference runs with USE_GPU=1).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
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


from dli_b200.measure import ClockSampler, decode_step_bytes, prefill_flops  # noqa: E402


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

    # ---- decode through the plugin call (POST /inference on the drop-in worker) ---------------
    if B == 1:
        e2e = e2e_through_worker(eng, prompt[0], ctx, K)
    else:  # the worker serves one prompt per request: larger batches go through the per-step host call
        e2e = e2e_step_host(eng, B, K)

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
    extra["env"] = {k: v for k, v in os.environ.items() if k.startswith("DLI_")}  # knobs that change the measured path
    if args.batch == 1 and not args.no_extra:
        try:
            extra["decode_b32"] = batch_decode_extra(spec, 32, ctx, min(K, 32), peaks, args.page_size, eng)
        except Exception as e:  # noqa: BLE001
            extra["decode_b32"] = {"error": str(e)[:200]}
    if not args.no_extra:
        eng_w, eng_layers = eng.w, eng.layers
        try:
            eng.w, eng.layers = {}, []
            eng.unload()  # frees KV and workspaces; the weights go with the references below
            del eng_w, eng_layers
            torch.cuda.empty_cache()
            extra["hf_gpu"] = hf_gpu_comparator(spec, ctx)
            extra["hf_gpu"]["ours_over_hf_decode"] = value / extra["hf_gpu"]["decode_tokens_per_s"]
        except Exception as e:  # noqa: BLE001
            extra["hf_gpu"] = {"error": str(e)[:300]}
    return {"value": value, "ms_per_step": ms_step, "e2e": e2e, "roofline": roof, "gpu_launches": launches,
            "clocks": clk.summary(), "extra": extra,
            "config": {"workload": f"{args.model} bf16 random-init, 1 GPU, batch {B} greedy decode at {ctx}-token context, "
                                   f"paged KV (page {args.page_size}); BASELINE.json configs[2]",
                       "batch": B, "ctx": ctx, "page_size": args.page_size, "parallelism": "1 stage",
                       "l2": f"inputs larger than L2: {step_bytes / 1e9:.2f} GB streamed per step vs 126 MB L2"}}


def e2e_step_host(eng, B, K):
    """Batch > 1: one decode step per call with HOST buffers (pinned tokens in, tokens out, sync per step)."""
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
    return {"value": B * K / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": 4 * B, "d2h_bytes_per_step": 4 * B,
            "ms_per_step": e2e_ms / K, "api": "ShardEngine.decode_step_host"}


def e2e_through_worker(eng, prompt_row, ctx, K):
    """The reference-facing plugin call: the drop-in Worker's POST /inference (worker.py, the route the
    reference master posts to: master/dashboard/views.py:345-355, :412-421), greedy.  The engine already
    holds the weights (the load routes are not part of the per-request path); everything a request does
    — JSON, tokenising the prompt text, the host->device copy of the ids, prefill, K decode steps, the
    device->host copy of the tokens, decoding to text — is inside the timed region."""
    from werkzeug.test import Client

    from dli_b200.worker import IdTokenizer, Worker

    w = Worker(env={"MODEL_CACHE_DIR": "/nonexistent", "USE_GPU": "1", "DLI_MAX_SEQ_LEN": str(eng.max_seq_len)})
    w.loaded_models["bench"] = eng
    w.loaded_tokenizers["bench"] = IdTokenizer()
    client = Client(w)
    text = " ".join(str(int(t)) for t in prompt_row)

    def request(n_new):
        t0 = time.perf_counter()
        r = client.post("/inference", json={"model_name": "bench", "prompt": text, "max_length": ctx + n_new,
                                            "do_sample": False, "timeout": 600})
        dt = time.perf_counter() - t0
        body = r.get_json()
        if r.status_code != 200 or body.get("status") != "success":
            raise RuntimeError(f"/inference failed: {r.status_code} {body}")
        n_out = len(body["result"].split())
        if n_out != ctx + n_new:
            raise RuntimeError(f"/inference returned {n_out} tokens, expected {ctx + n_new}")
        return dt

    request(1), request(K + 1)  # warm-up: graph capture for this state, allocator
    t_first = min(request(1) for _ in range(3))
    t_full = min(request(K + 1) for _ in range(2))
    dec = max(t_full - t_first, 1e-9)
    return {"value": K / dec, "unit": UNIT, "h2d_bytes_per_step": 4.0 * ctx / K, "d2h_bytes_per_step": 4.0 * (K + 1) / K,
            "ms_per_step": dec / K * 1e3, "api": "Worker.inference (POST /inference, werkzeug test client)",
            "request_s": t_full, "request_new_tokens": K + 1, "request_tokens_per_s": (K + 1) / t_full,
            "prefill_request_s": t_first, "prompt_tokens": ctx,
            "note": "value = K decode steps / (request with K+1 new tokens - request with 1 new token), both "
                    "timed on the host clock around the HTTP handler; h2d = the int32 prompt ids amortised over K"}


def hf_gpu_comparator(spec, ctx, n_new=32):
    """transformers generate (bf16, sdpa, greedy) on this GPU at the same shape: the stock-PyTorch GPU
    comparator BASELINE.md 5 names — what the reference's worker runs with USE_GPU=1
    (reference worker/app.py:23-26,121-124,297-305) on a torch that supports the B200."""
    import gc

    import transformers

    kw = dict(vocab_size=spec.vocab, hidden_size=spec.d_model, intermediate_size=spec.d_ff,
              num_hidden_layers=spec.n_layers, num_attention_heads=spec.n_heads, num_key_value_heads=spec.n_kv_heads,
              head_dim=spec.head_dim, max_position_embeddings=max(spec.max_pos, ctx + n_new + 8), rms_norm_eps=spec.eps,
              tie_word_embeddings=spec.tie_embeddings, rope_parameters={"rope_type": "default", "rope_theta": spec.rope_theta})
    cfg = (transformers.MistralConfig(sliding_window=spec.sliding_window, **kw) if spec.sliding_window
           else transformers.LlamaConfig(**kw))
    with torch.device("cuda"):
        m = transformers.AutoModelForCausalLM.from_config(cfg, dtype=torch.bfloat16).eval()
    ids = torch.randint(0, spec.vocab, (1, ctx), generator=torch.Generator().manual_seed(1)).cuda()
    mask = torch.ones_like(ids)

    def gen(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with torch.no_grad():
            m.generate(ids, attention_mask=mask, max_new_tokens=n, min_new_tokens=n, do_sample=False, pad_token_id=0)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    gen(2)
    t1 = min(gen(1) for _ in range(2))
    t2 = min(gen(1 + n_new) for _ in range(2))
    out = {"decode_tokens_per_s": n_new / max(t2 - t1, 1e-9), "prefill_tokens_per_s": ctx / t1, "ms_per_step": (t2 - t1) / n_new * 1e3,
           "impl": f"transformers {transformers.__version__} generate, bf16, attn_implementation={m.config._attn_implementation}, "
                   f"torch {torch.__version__}", "sample": f"batch 1, {ctx}-token prompt, {n_new} decode steps"}
    del m
    gc.collect()
    torch.cuda.empty_cache()
    return out


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
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this kernel from an `ncu --set full` capture
    # (profiles/): a recorded measurement, not taken in this run — the capture's tree is named beside it
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        rec = json.load(open(tp))
        traffic = rec.get("gate_up_projection_dram_bytes_per_launch")
        traffic_src = rec.get("source")
    return {"bound": "hbm", "kernel": "gemv_kernel<M,gated> (gate/up projection)" if B <= 4 else
            "gemm_tcgen05_kernel<swap,gated> (gate/up projection)", "achieved": ach, "peak": peaks["hbm_gbs"],
            "unit": "GB/s", "frac": ach / peaks["hbm_gbs"], "traffic": traffic, "traffic_source": traffic_src,
            "bytes_per_launch": nbytes,
            "us_per_launch": ms * 1e3, "peak_source": peaks["source"]}


def batch_decode_extra(spec, B, ctx, K, peaks, page_size, old_eng):
    """Extra (not the headline): batch-32 decode at the same context."""
    from dli_b200.engine import ShardEngine

    max_len = ctx + 3 * K + 16
    eng = ShardEngine(spec, max_batch=B, max_seq_len=max_len, page_size=page_size)
    eng.w, eng.layers = old_eng.w, old_eng.layers  # share the weights already in HBM
    g = torch.Generator().manual_seed(2)
    ids = torch.randint(0, spec.vocab, (B * ctx,), generator=g, dtype=torch.int32).cuda()
    eng.start_batch(B, ctx, max_len)
    pf_ms = ev_time_ms(lambda: eng.prefill(B, ctx, ids=ids))
    for _ in range(3):
        eng.decode_step(B)
    runs = [ev_time_ms(lambda: [eng.decode_step(B) for _ in range(K)]) / K for _ in range(3)]
    ms = statistics.median(runs)
    nbytes = decode_step_bytes(spec, spec.n_layers, True, B, ctx + 3 + 3 * K // 2)
    out = {"tokens_per_s": B / (ms / 1e3), "ms_per_step": ms, "ms_per_step_runs": runs,
           "hbm_frac": nbytes / (ms / 1e3) / 1e9 / peaks["hbm_gbs"],
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
    out = {"value": r["decode_tok_s"], "unit": UNIT, "cores": r["threads"], "kind": "port",
           "sample": r["sample"] + f" (transformers generate, the reference's worker/app.py path with USE_GPU=0); "
                                   f"host cpu_count={os.cpu_count()}", "wall_s": time.time() - t0,
           "prefill_tok_s": r["prefill_tok_s"]}
    try:
        out["configs0"] = cpu_configs0()
    except Exception as e:  # noqa: BLE001
        out["configs0"] = {"error": str(e)[:200]}
    return out


def cpu_configs0():
    """BASELINE.json configs[0]: GPT-2 small (124M) random-init, fp32 (the reference's default load dtype,
    worker/app.py:121), USE_GPU=0, 32-token prompt -> 64 new tokens (max_length 96) through the port of the
    reference's /inference body — greedy (the parity mode) and with the reference's own sampling call."""
    from dli_b200.engine import NAMED_SPECS
    from oracle import ref_worker_port as P

    spec = NAMED_SPECS["gpt2"]
    model = P.build_cpu_model(spec, dtype=torch.float32)
    greedy = P.time_cpu_decode(spec, 32, 64, dtype=torch.float32, greedy=True, model=model, warmup_tokens=1)
    sampled = P.time_cpu_decode(spec, 32, 64, dtype=torch.float32, greedy=False, model=model, warmup_tokens=0)
    return {"workload": "GPT-2 small fp32 random-init, 1 worker, 1 shard, USE_GPU=0, 32-token prompt -> 64 tokens",
            "cores": greedy["threads"], "kind": "port",
            "greedy": {"request_s": greedy["t_total_s"], "tokens_per_s": 64 / greedy["t_total_s"],
                       "decode_tokens_per_s": greedy["decode_tok_s"]},
            "reference_sampling": {"request_s": sampled["t_total_s"], "tokens_per_s": 64 / sampled["t_total_s"],
                                   "decode_tokens_per_s": sampled["decode_tok_s"]}}


def run_reference(args, spec):
    """--impl reference: the reference's CPU path (port, see oracle/ref_worker_port.py),
    all host threads, each step = one decoded token at the product arm's context (batch 1, `--ctx` tokens), decode
    rate = K tokens / (t[K+1 new tokens] - t[1 new token])."""
    from oracle import ref_worker_port as P

    _use_all_host_threads()
    model = P.build_cpu_model(spec)
    tok = P.IdTokenizer()
    K, W = min(args.steps, 64), min(args.warmup, 2)  # bounded: ~0.12 s per token on the box's 16 cores
    # the product arm's context (DLI_REF_CTX bounds it: the prompt pass of an 8B model costs the host cores ~10-20 s
    # per call and the arm makes two of them)
    prompt_len = max(1, min(int(os.environ.get("DLI_REF_CTX", args.ctx)), spec.max_pos - K - 2))
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, spec.vocab, (prompt_len,), generator=g)
    prompt = " ".join(str(int(i)) for i in ids)
    if W:  # warm-up on a short prompt: thread pool, allocator, first-call costs
        short = " ".join(str(int(i)) for i in ids[:16])
        P.run_inference_port(model, tok, short, min(16, prompt_len) + 1 + W, timeout=1e9, greedy=True)
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
            "config": {"workload": f"{args.model} bf16 random-init, batch 1 greedy decode at {prompt_len}-token context "
                                   f"(CPU, {K} steps)", "batch": 1, "ctx": prompt_len},
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

        import torch.distributed as dist

        try:
            res = run_pipeline_bench(args, spec, peaks)
        finally:
            if dist.is_initialized():
                dist.destroy_process_group()
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
