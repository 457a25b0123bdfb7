#!/bin/bash
# Quick GPU iteration: parity tests of the path, then the c2 and c3 (batch 8) bench lines.  Run under gpurun.
python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py -m gpu -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_quick.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
python bench.py --workload c3 --batch 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3.log 2>&1
tail -2 gpurun_out/pytest_quick.log
python tools/show_bench.py bench_c2 bench_c3
