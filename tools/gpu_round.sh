#!/bin/bash
# Round-end evidence run (under gpurun, one GPU): widened-op timings, the remaining bench lines, then the ncu launch list.
python -m pytest tests/test_gpu_widened_speed.py -m gpu -q > gpurun_out/widened.log 2>&1
python bench.py --workload c1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c1.log 2>&1
python bench.py --workload c3 --batch 64 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_b64.log 2>&1
python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1
python bench.py --workload c5 --batch 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_b1.log 2>&1
python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/bench_c2_eager.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
    python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
