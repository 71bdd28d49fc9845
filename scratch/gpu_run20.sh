#!/bin/bash
# round 2, GPU call 20 (1 GPU): whole GPU suite and the bench after the prefill-attention rewrite
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/t20_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t20_tests.log
tail -4 gpurun_out/t20_tests.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/t20_bench.log 2> gpurun_out/t20_bench.err; echo "bench rc=$?"
python - <<PY
import json
for ln in open('gpurun_out/t20_bench.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('value',round(d['value'],1),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value'],1),'prefill',round(e['prefill_tokens_per_s']),round(e['prefill_roofline']['frac'],4),'b32',round(e['decode_b32']['tokens_per_s']),round(e['decode_b32']['hbm_frac'],3), 'b32 prefill', round(e['decode_b32']['prefill_tokens_per_s']), 'hf', e.get('hf_gpu',{}).get('decode_tokens_per_s'))
PY
