#!/bin/bash
# round 2, GPU call 14 (1 GPU): prefill attention with one-instruction exp2, lazy reference maximum, one TMEM wait per
# tile, per-quadrant pair barriers — kernel tests, attention time alone, full-depth parity, short bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_shapes_gpu.py -m gpu -q -x > gpurun_out/t14_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t14_tests.log
tail -3 gpurun_out/t14_tests.log | cut -c1-300
timeout 200 python scratch/exp_attn_prefill_time.py > gpurun_out/t14_attn_time.log 2>&1; cat gpurun_out/t14_attn_time.log | tail -3
timeout 900 python -m pytest tests/test_full_depth_gpu.py -m gpu -q -x -s > gpurun_out/t14_full_depth.log 2>&1; echo "rc=$?" >> gpurun_out/t14_full_depth.log
grep -v "^$" gpurun_out/t14_full_depth.log | tail -8 | cut -c1-400
timeout 600 python bench.py --steps 32 --warmup 4 > gpurun_out/t14_bench.log 2> gpurun_out/t14_bench.err; echo "bench rc=$?"
python - <<PY
import json
for ln in open('gpurun_out/t14_bench.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('value',round(d['value'],1),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value'],1),'prefill',round(e['prefill_tokens_per_s']),round(e['prefill_roofline']['frac'],4),'b32',round(e['decode_b32']['tokens_per_s']),round(e['decode_b32']['hbm_frac'],3), 'b32 prefill', round(e['decode_b32']['prefill_tokens_per_s']))
PY
