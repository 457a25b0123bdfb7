#!/bin/bash
# Standard GPU check: parity tests, then the c2 and (batch 8) c3 bench lines.  Run under gpurun from the repo root.
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
python bench.py --workload c3 --batch 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3.log 2>&1
python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.log 2>&1
