#!/bin/bash
# First-pass candidate target of the soft-mask tiles (DIBR_SOFT_SEL) with the distribution sort in place.
NAMES=""
run() { tag=$1; shift; for w in c3:8 c3:64 c5: c2:; do wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb};
  env "$@" timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/sel_${tag}_$wl$nb.log 2>&1; NAMES="$NAMES sel_${tag}_$wl$nb"; done; }
run d DIBR_X=0
run 128 DIBR_B200_TUNING=1 DIBR_SOFT_SEL=128
run 512 DIBR_B200_TUNING=1 DIBR_SOFT_SEL=512
run 1024 DIBR_B200_TUNING=1 DIBR_SOFT_SEL=1024
python tools/show_bench.py $NAMES
