#!/bin/bash
# round 2, GPU call 7 (1 GPU): stream-K weight stream tests + A/B at batch 32, one-graph ticks in mailbox mode
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_mailbox_gpu.py tests/test_engine_gpu.py tests/test_batching_gpu.py tests/test_shapes_gpu.py -m gpu -q -x -k "splitk or gemm or mailbox or batched or batching or shapes" > gpurun_out/t7_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t7_tests.log
tail -25 gpurun_out/t7_tests.log
for k in 0 1; do
  DLI_STREAMK=$k timeout 300 python bench.py --no-cpu-baseline --steps 32 > gpurun_out/t7_bench_sk$k.log 2> gpurun_out/t7_bench_sk$k.err
  python - <<PY
import json
for ln in open('gpurun_out/t7_bench_sk$k.log'):
    if ln.startswith('{'):
        d=json.loads(ln); b=d['extra'].get('decode_b32',{})
        print('streamk=$k', 'b1', round(d['value'],1), 'b32', b.get('tokens_per_s'), b.get('ms_per_step_runs'), b.get('hbm_frac'), 'prefill', round(d['extra']['prefill_tokens_per_s']))
PY
done
tail -3 gpurun_out/t7_bench_sk1.err
