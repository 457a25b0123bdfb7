#!/bin/bash
# Compare builds of the library (DIBR_B200_LIB=build/*.so) and env knobs on the c2 / c3 bench lines.  Run under gpurun.
#   bash tools/gpu_variants.sh "ENV1=.. ENV2=.." "ENV=.." ...      (each argument = one variant's environment)
i=0
for v in "$@"; do
  i=$((i+1))
  env $v python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/var${i}_c2.log 2>&1
  env $v python bench.py --workload c3 --batch 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/var${i}_c3.log 2>&1
  echo "== variant $i: $v"; python tools/show_bench.py var${i}_c2 var${i}_c3
done
