#!/bin/bash
# Last check of the final tree: parity suite, smoke, default bench (with reference arm), per-workload lines.
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
for w in c3:8 c2: c5: c1: c4:16 c4:; do wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb};
  timeout 300 python bench.py --workload $wl $B --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_$wl${nb:+_b$nb}.log 2>&1; done
python tools/show_bench.py bench_default bench_c3_b8 bench_c2 bench_c5 bench_c1 bench_c4_b16 bench_c4
