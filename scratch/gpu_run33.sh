#!/bin/bash
# round 2, GPU call 33 (4 GPUs): the 4-stage bench on the final tree (parity check inside), Mistral named configuration
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 4 > gpurun_out/t33_bench4.log 2> gpurun_out/t33_bench4.err; echo "bench4 rc=$?"
python - <<PY
import json
for ln in open('gpurun_out/t33_bench4.log'):
    if ln.startswith('{'):
        d=json.loads(ln); e=d['extra']
        print('value',round(d['value'],1),'tick',round(e['tick_ms'],4),'frac',round(d['roofline']['frac'],3),'parity',e['parity_check'],'e2e',round(d['e2e']['value'],1))
        nc=e.get('named_config')
        if nc: print('   named', {k:(round(v,3) if isinstance(v,float) else v) for k,v in nc.items() if k not in ('note','workload','clocks')})
PY
grep -v "unbatched P2P\|OMP_NUM\|\*\*\*\*\|^$" gpurun_out/t33_bench4.err | tail -5
