#!/bin/bash
# Round 2, last check of the final tree on one GPU: what the driver runs + the per-workload bench lines.
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1
timeout 900 python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_c3_b8.log 2>&1
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
timeout 300 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1
timeout 300 python bench.py --workload c1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.log 2>&1
timeout 300 python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b16.log 2>&1
timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b128.log 2>&1
python tools/show_bench.py bench_default bench_c3_b8 bench_c2 bench_c5 bench_c1 bench_c4_b16 bench_c4_b128
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.log').read().strip().splitlines()[-1])
print("c4 extra:", {k: d['c4'][k] for k in ('ms_per_step','all_reduce_ms')}, d['c4']['e2e']['ms_per_step'], "roofline", d['roofline']['kernel'], round(d['roofline']['frac'],4))
PY
