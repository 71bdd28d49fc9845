#!/bin/bash
# round 2, GPU call 6 (8 GPUs): NCCL pipeline tests (2 and 4 stages, worker group), bench at 8 / 4 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline_gpu.py -m gpu -q --durations=5 > gpurun_out/t6_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t6_tests.log
tail -12 gpurun_out/t6_tests.log
run() { n=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $n "$@"; }
run 8 > gpurun_out/t6_bench8.log 2> gpurun_out/t6_bench8.err; tail -c 4500 gpurun_out/t6_bench8.log; grep -v "unbatched P2P\|OMP_NUM\|\*\*\*\*" gpurun_out/t6_bench8.err | tail -8
DLI_PIPE_HOP=1 run 8 --no-extra > gpurun_out/t6_bench8_hop1.log 2> gpurun_out/t6_bench8_hop1.err; head -c 300 gpurun_out/t6_bench8_hop1.log; echo; grep -o '"tick_ms": [0-9.]*' gpurun_out/t6_bench8_hop1.log | head -2
run 4 > gpurun_out/t6_bench4.log 2> gpurun_out/t6_bench4.err; tail -c 3000 gpurun_out/t6_bench4.log; grep -v "unbatched P2P\|OMP_NUM\|\*\*\*\*" gpurun_out/t6_bench4.err | tail -5
