#!/bin/bash
# round 2, GPU call 13 (8 GPUs): final tree — NCCL tests, bench at 8 GPUs (70B named configuration with the head on
# the last stage), bench at 2 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline_gpu.py tests/test_two_devices_gpu.py -m gpu -q > gpurun_out/t13_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t13_tests.log
tail -4 gpurun_out/t13_tests.log | cut -c1-300
run() { n=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $n "$@"; }
run 8 > gpurun_out/t13_bench8.log 2> gpurun_out/t13_bench8.err
run 2 > gpurun_out/t13_bench2.log 2> gpurun_out/t13_bench2.err
python - <<PY
import json
for f in ("t13_bench8","t13_bench2"):
    for ln in open(f'gpurun_out/{f}.log'):
        if ln.startswith('{'):
            d=json.loads(ln); e=d['extra']
            print(f, 'value',round(d['value'],1),'tick',round(e['tick_ms'],4),'frac',round(d['roofline']['frac'],3),'parity',e['parity_check'],'e2e',round(d['e2e']['value'],1))
            nc=e.get('named_config')
            if nc: print('   named', {k:(round(v,3) if isinstance(v,float) else v) for k,v in nc.items() if k not in ('note','workload','clocks')})
PY
grep -v "unbatched P2P\|OMP_NUM\|\*\*\*\*\|^$" gpurun_out/t13_bench8.err | tail -5
