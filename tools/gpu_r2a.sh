#!/bin/bash
# Round 2, first GPU call: smoke, the whole GPU suite (no -x: see every failure), the default bench line (c3, 64 views
# + the c4 key), the reference arm, and ncu --set full captures of one c3 (64 views) and one c5 step.
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1
timeout 300 python bench.py --workload c3 --batch 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_b8.log 2>&1
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1
timeout 600 bash tools/prof_all.sh c3 64
timeout 600 bash tools/prof_all.sh c5
tail -5 gpurun_out/pytest_gpu.log
python tools/show_bench.py bench_default bench_c3_b8 bench_c2 bench_c5
