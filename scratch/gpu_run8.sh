#!/bin/bash
# round 2, GPU call 8 (1 GPU): 240-column prefill tiles, one-graph mailbox ticks, full GPU test suite, profiles
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/t8_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t8_tests.log
tail -14 gpurun_out/t8_tests.log
timeout 400 python bench.py > gpurun_out/t8_bench.log 2> gpurun_out/t8_bench.err
python - <<PY
import json
for ln in open('gpurun_out/t8_bench.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('b1', round(d['value'],1), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['frac'],3), 'step', round(e['roofline_step']['frac'],4),
              'prefill', round(e['prefill_tokens_per_s']), round(e['prefill_roofline']['frac'],4), 'b32', e['decode_b32'].get('tokens_per_s'), e['decode_b32'].get('hbm_frac'), 'hf', e['hf_gpu'].get('decode_tokens_per_s'))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t8_ncu1.log 2>&1
gzip -f gpurun_out/r2_launches.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 300 -c 8 -o gpurun_out/r2_gemv python bench.py --steps 4 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t8_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attn_decode -s 100 -c 2 -o gpurun_out/r2_attn python bench.py --steps 4 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t8_ncu3.log 2>&1
ls -la gpurun_out | grep -E "r2_|t8_"
