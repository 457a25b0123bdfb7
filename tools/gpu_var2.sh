#!/bin/bash
# Variant libraries on c3-8 / c3-64 / c5 only (quick).
NAMES=""
for v in $VARIANTS; do
  L=""; [ "$v" != "main" ] && L="$PWD/build/variants/$v.so"
  for w in c3:8 c3:64 c5:; do
    wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb}
    DIBR_B200_LIB=$L timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/v_${v}_$wl$nb.log 2>&1
    NAMES="$NAMES v_${v}_$wl$nb"
  done
done
python tools/show_bench.py $NAMES
