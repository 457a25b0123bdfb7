#!/bin/bash
# Round 2: parity suite (with the from-mesh / loss-step tests) + default bench (c3-64 + c4 extra) + c4 at 16 cameras.
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^E  |^FAILED|^ERROR" gpurun_out/pytest_gpu.log | sort | uniq -c | sort -rn | head -12
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
timeout 300 python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b16.log 2>&1
timeout 300 python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/bench_c4_b16_eager.log 2>&1
timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_c3_b8.log 2>&1
python tools/show_bench.py bench_default bench_c4_b16 bench_c4_b16_eager bench_c3_b8
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.log').read().strip().splitlines()[-1])
print("c4 extra:", {k: d['c4'][k] for k in ('ms_per_step','all_reduce_ms','launch_mode')}, d['c4']['e2e']['ms_per_step'])
for n in ('bench_c4_b16','bench_c4_b16_eager'):
    d=json.loads(open(f'gpurun_out/{n}.log').read().strip().splitlines()[-1]); print(n, d['ms_per_step'], d['config'].get('launch_mode'))
PY
tail -3 gpurun_out/bench_default.err
