#!/bin/bash
# CTA-size knobs (DIBR_B200_TUNING=1) on c3-8 / c3-64 / c5 with the final kernels.
NAMES=""
run() { tag=$1; shift; for w in c3:8 c3:64 c5:; do wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb};
  env DIBR_B200_TUNING=1 "$@" timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/nt_${tag}_$wl$nb.log 2>&1; NAMES="$NAMES nt_${tag}_$wl$nb"; done; }
run base DIBR_X=0
run r128 DIBR_RASTER_NT=128
run s64 DIBR_SOFT_NT=64
run s256 DIBR_SOFT_NT=256
python tools/show_bench.py $NAMES
