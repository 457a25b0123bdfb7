#!/bin/bash
# Round 2: raster_bwd reduction-depth variants; ncu launch lists of the c4 step (16 and 128 cameras) and the default
# bench; refreshed c3 (8 views) per-line profile.
NAMES=""
timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_c3_b8.log 2>&1
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
for v in $VARIANTS; do
  DIBR_B200_LIB=$PWD/build/variants/$v.so timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/var_${v}_c3.log 2>&1
  DIBR_B200_LIB=$PWD/build/variants/$v.so timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/var_${v}_c2.log 2>&1
  NAMES="$NAMES var_${v}_c3 var_${v}_c2"
done
python tools/show_bench.py bench_c3_b8 bench_c2 $NAMES
for nb in 16 128; do
  timeout 300 python bench.py --workload c4 --batch $nb --steps 3 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/bench_c4_b$nb.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_c4_b$nb.csv \
      python bench.py --workload c4 --batch $nb --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/ncu_launches_c4_b$nb.log 2>&1
  tail -c 600 gpurun_out/bench_c4_b$nb.log
done
bash tools/prof_all.sh c3 8
