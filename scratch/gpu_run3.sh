#!/bin/bash
# round 2, GPU call 3 (2 GPUs): mailbox pipeline tests (1-GPU emulation + NCCL), pipeline bench hop 2 / hop 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_mailbox_gpu.py tests/test_pipeline_gpu.py tests/test_kernels_gpu.py -m gpu -q -k "mailbox or nccl or worker_group or cluster" --durations=5 > gpurun_out/t3_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t3_tests.log
tail -15 gpurun_out/t3_tests.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 "$@"; }
run --steps 32 --warmup 8 > gpurun_out/t3_bench2.log 2> gpurun_out/t3_bench2.err; tail -c 2500 gpurun_out/t3_bench2.log; tail -5 gpurun_out/t3_bench2.err
DLI_PIPE_HOP=1 run --steps 32 --warmup 8 --no-extra > gpurun_out/t3_bench2_hop1.log 2> gpurun_out/t3_bench2_hop1.err; tail -c 1200 gpurun_out/t3_bench2_hop1.log
timeout 300 python bench.py --no-cpu-baseline --no-extra > gpurun_out/t3_bench1.log 2>&1; tail -c 400 gpurun_out/t3_bench1.log
