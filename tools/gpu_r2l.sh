#!/bin/bash
# Round 2: ordered soft-mask work list on / off (DIBR_SOFT_WORKLIST), bench lines only.
for w in c3:8 c2: c5: c3:64; do
  wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb}
  timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/wl_on_$wl$nb.log 2>&1
  DIBR_B200_TUNING=1 DIBR_SOFT_WORKLIST=0 timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/wl_off_$wl$nb.log 2>&1
done
python tools/show_bench.py wl_on_c38 wl_off_c38 wl_on_c364 wl_off_c364 wl_on_c2 wl_off_c2 wl_on_c5 wl_off_c5
