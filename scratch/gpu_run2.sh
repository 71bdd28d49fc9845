#!/bin/bash
# round 2, GPU call 2: all GPU tests, GEMV dealing A/B, ncu launch lists of the decode step
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s --durations=10 > gpurun_out/t2_tests.log 2>&1; echo "rc=$?" >> gpurun_out/t2_tests.log
tail -8 gpurun_out/t2_tests.log
for d in 0 1 2; do
  DLI_GEMV_DEAL=$d timeout 200 python bench.py --no-extra --no-cpu-baseline > gpurun_out/t2_bench_deal$d.log 2>&1
  tail -c 300 gpurun_out/t2_bench_deal$d.log | head -c 300; echo
done
DLI_GEMV_DEAL=0 DLI_ATTN_WIDE=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
  --log-file gpurun_out/t2_launches_old.csv python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t2_ncu_old.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
  --log-file gpurun_out/t2_launches_new.csv python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/t2_ncu_new.log 2>&1
gzip -f gpurun_out/t2_launches_old.csv gpurun_out/t2_launches_new.csv
ls -la gpurun_out
