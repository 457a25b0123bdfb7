#!/bin/bash
# Round 2: parity suite + the c4 step (16 / 128 cameras) after the per-element Adam.
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^E  |^FAILED|^ERROR" gpurun_out/pytest_gpu.log | sort | uniq -c | sort -rn | head -12
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b16.log 2>&1
timeout 300 python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/bench_c4_b16_eager.log 2>&1
timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b128.log 2>&1
python tools/show_bench.py bench_c4_b16 bench_c4_b16_eager bench_c4_b128
