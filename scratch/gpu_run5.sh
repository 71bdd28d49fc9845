#!/bin/bash
# round 2, GPU call 5 (1 GPU): GEMV tests, pipelined GEMV A/B in the decode step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_mailbox_gpu.py -m gpu -q -k "gemv or lm_head or engine or mailbox or graph" > gpurun_out/t5_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t5_tests.log
tail -6 gpurun_out/t5_tests.log
for p in 0 8; do
  DLI_GEMV_PIPE=$p timeout 200 python bench.py --no-extra --no-cpu-baseline > gpurun_out/t5_bench_pipe$p.log 2>&1
  python - <<PY
import json
for ln in open('gpurun_out/t5_bench_pipe$p.log'):
    if ln.startswith('{'):
        d=json.loads(ln); print('pipe$p', round(d['value'],1), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], round(d['extra']['roofline_step']['frac'],4))
PY
done
