#!/bin/bash
# ncu --set full capture of the four hot kernels for one step of a workload (run under gpurun).
#   bash tools/prof_all.sh c2 [batch]
W=${1:-c2}; B=${2:+--batch $2}
python tools/prof_step.py --workload $W $B > gpurun_out/prof_plain_$W.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:raster|soft_mask" -f \
    -o gpurun_out/prof_$W python tools/prof_step.py --workload $W $B > gpurun_out/prof_ncu_$W.log 2>&1
