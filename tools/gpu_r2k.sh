#!/bin/bash
# Round 2: ordered soft-mask work list on / off (DIBR_SOFT_WORKLIST) — parity suite + bench lines both ways.
timeout 1500 python -m pytest tests -m gpu -q -x ${PYTEST_K} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^E  |^FAILED|^ERROR" gpurun_out/pytest_gpu.log | sort | uniq -c | sort -rn | head -12
tail -3 gpurun_out/pytest_gpu.log
for w in c3:8 c3:64 c2: c5: c4:16 c1:; do
  wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb}
  timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/wl_on_$wl$nb.log 2>&1
  DIBR_B200_TUNING=1 DIBR_SOFT_WORKLIST=0 timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/wl_off_$wl$nb.log 2>&1
done
python tools/show_bench.py wl_on_c38 wl_off_c38 wl_on_c364 wl_off_c364 wl_on_c2 wl_off_c2 wl_on_c5 wl_off_c5 wl_on_c416 wl_off_c416 wl_on_c1 wl_off_c1
