#!/bin/bash
# round 2, GPU call 30 (1 GPU): final tree — whole GPU suite, bench (both arms), ncu launch list and full captures
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/t30_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t30_tests.log
tail -3 gpurun_out/t30_tests.log | cut -c1-200
timeout 600 python bench.py > gpurun_out/t30_bench.log 2> gpurun_out/t30_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/t30_bench_ref.log 2> gpurun_out/t30_bench_ref.err; echo "ref rc=$?"
python - <<PY
import json
for f in ("t30_bench","t30_bench_ref"):
    for ln in open(f'gpurun_out/{f}.log'):
        if ln.startswith('{'):
            d=json.loads(ln); e=d.get('extra',{})
            if d.get('impl')=='reference': print('reference', round(d['value'],2), d['unit']); continue
            print('b1', round(d['value'],1), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['frac'],3), 'step', round(e['roofline_step']['frac'],4),
                  'prefill', round(e['prefill_tokens_per_s']), round(e['prefill_roofline']['frac'],4), 'b32', e['decode_b32'].get('tokens_per_s'), e['decode_b32'].get('hbm_frac'), 'hf', e['hf_gpu'].get('decode_tokens_per_s'), d['clocks'])
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2b_launches.csv python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t30_ncu1.log 2>&1
gzip -f gpurun_out/r2b_launches.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 300 -c 6 -o gpurun_out/r2b_gemv python bench.py --steps 4 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t30_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:attn_prefill_tc|gemm_prefill|rmsnorm_rows" -s 40 -c 9 -o gpurun_out/r2b_prefill python scratch/prefill_only.py 2 > gpurun_out/t30_ncu3.log 2>&1
ls -la gpurun_out | grep -E "r2b_|t30_"
