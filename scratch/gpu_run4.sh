#!/bin/bash
# round 2, GPU call 4 (2 GPUs): worker-group serving test, pipeline bench with the request-path e2e
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pipeline_gpu.py -m gpu -q -x -k "worker_group" --durations=3 > gpurun_out/t4_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t4_tests.log
tail -30 gpurun_out/t4_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 32 --warmup 8 > gpurun_out/t4_bench2.log 2> gpurun_out/t4_bench2.err
tail -c 3500 gpurun_out/t4_bench2.log; tail -8 gpurun_out/t4_bench2.err
