#!/bin/bash
# round 2, GPU call 31 (2 GPUs): NCCL tests and the 2-stage bench on the final tree
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline_gpu.py tests/test_two_devices_gpu.py -m gpu -q > gpurun_out/t31_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t31_tests.log
tail -4 gpurun_out/t31_tests.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 > gpurun_out/t31_bench2.log 2> gpurun_out/t31_bench2.err; echo "bench2 rc=$?"
python - <<PY
import json
for ln in open('gpurun_out/t31_bench2.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('value',round(d['value'],1),'tick',round(e['tick_ms'],4),'frac',round(d['roofline']['frac'],3),'parity',e['parity_check'],'e2e',round(d['e2e']['value'],1))
        nc=e.get('named_config')
        if nc: print('   named', {k:(round(v,3) if isinstance(v,float) else v) for k,v in nc.items() if k not in ('note','workload','clocks')})
PY
grep -v "unbatched P2P\|OMP_NUM\|\*\*\*\*\|^$" gpurun_out/t31_bench2.err | tail -5
