#!/bin/bash
# Soft-mask CTA size: density heuristic (default) vs forced 128 / 256 on c4, c2, c1, c5; then the parity suite.
NAMES=""
run() { tag=$1; shift; for w in c4:16 c4:128 c2: c1: c5:; do wl=${w%%:*}; nb=${w##*:}; B=${nb:+--batch $nb};
  env "$@" timeout 300 python bench.py --workload $wl $B --steps 8 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/nt_${tag}_$wl$nb.log 2>&1; NAMES="$NAMES nt_${tag}_$wl$nb"; done; }
run auto DIBR_X=0
run s128 DIBR_B200_TUNING=1 DIBR_SOFT_NT=128
run s256 DIBR_B200_TUNING=1 DIBR_SOFT_NT=256
python tools/show_bench.py $NAMES
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
