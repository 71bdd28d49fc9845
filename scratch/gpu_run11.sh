#!/bin/bash
# round 2, GPU call 11 (1 GPU): what the driver runs at round end — all GPU tests, smoke(), bench (both arms), timed
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/t11_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t11_tests.log
tail -8 gpurun_out/t11_tests.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) 2>&1 | tail -5
( time timeout 900 python bench.py --impl reference ) > gpurun_out/t11_ref.log 2>&1; tail -c 900 gpurun_out/t11_ref.log
( time timeout 900 python bench.py ) > gpurun_out/t11_bench.log 2> gpurun_out/t11_bench.err; tail -4 gpurun_out/t11_bench.err
python - <<PY
import json
for ln in open('gpurun_out/t11_bench.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('b1', round(d['value'],1), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['frac'],3), 'traffic', d['roofline']['traffic'], 'step', round(e['roofline_step']['frac'],4),
              'prefill', round(e['prefill_tokens_per_s']), round(e['prefill_roofline']['frac'],4), 'b32', round(e['decode_b32']['tokens_per_s']), round(e['decode_b32']['hbm_frac'],3), 'hf', round(e['hf_gpu']['decode_tokens_per_s'],1), 'cpu', round(d['cpu_baseline']['value'],2), d['cpu_baseline']['configs0']['greedy'], 'launches', d['gpu_launches'], d['clocks'])
PY
