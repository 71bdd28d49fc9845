#!/bin/bash
# round 2, GPU call 9 (2 GPUs): worker group with the vocabulary-parallel head, two-device test, pipeline bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline_gpu.py tests/test_two_devices_gpu.py tests/test_mailbox_gpu.py -m gpu -q -x --durations=3 > gpurun_out/t9_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t9_tests.log
tail -40 gpurun_out/t9_tests.log | cut -c1-400
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 32 --warmup 8 > gpurun_out/t9_bench2.log 2> gpurun_out/t9_bench2.err
python - <<PY
import json
for ln in open('gpurun_out/t9_bench2.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('N=2 value', round(d['value'],1), 'tick', round(e['tick_ms'],4), 'parity', e['parity_check'])
        print(' e2e', {k:(round(v,2) if isinstance(v,float) else v) for k,v in d['e2e'].items() if k not in ('note','api')})
PY
grep -v "unbatched P2P\|OMP_NUM\|\*\*\*\*\|^$" gpurun_out/t9_bench2.err | tail -6
