#!/bin/bash
# Round 2: parity suite + bench lines + variant libraries (DIBR_B200_LIB) on c3 (8 views) and c2.
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_c3_b8.log 2>&1
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1
NAMES="bench_default bench_c3_b8 bench_c2 bench_c5"
for v in $VARIANTS; do
  DIBR_B200_LIB=$PWD/build/variants/$v.so timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/var_${v}_c3.log 2>&1
  DIBR_B200_LIB=$PWD/build/variants/$v.so timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/var_${v}_c2.log 2>&1
  NAMES="$NAMES var_${v}_c3 var_${v}_c2"
done
grep -E "^E  |^FAILED|^ERROR" gpurun_out/pytest_gpu.log | sort | uniq -c | sort -rn | head -12
tail -4 gpurun_out/pytest_gpu.log
python tools/show_bench.py $NAMES
