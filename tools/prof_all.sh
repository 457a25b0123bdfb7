#!/bin/bash
# ncu --set full capture of the hot kernels (binning included) for one step of a workload (run under gpurun).
# The report is turned into CSVs on the box (raw page + per-instruction source page, gzipped); the .ncu-rep itself is
# kept only when small (gpurun_out/ is limited to 64 MiB).
#   bash tools/prof_all.sh c2 [batch] [fused]
W=${1:-c2}; B=${2:+--batch $2}; FUSED=${3:-1}
TAG=$W${2:+_b$2}
python tools/prof_step.py --workload $W $B --fused $FUSED > gpurun_out/prof_plain_$TAG.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k "regex:raster|soft_mask|setup_count|fill_kernel" -f \
    -o /tmp/prof_$TAG python tools/prof_step.py --workload $W $B --fused $FUSED > gpurun_out/prof_ncu_$TAG.log 2>&1
ncu -i /tmp/prof_$TAG.ncu-rep --page raw --csv 2>/dev/null | gzip > gpurun_out/prof_${TAG}_raw.csv.gz
ncu -i /tmp/prof_$TAG.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip > gpurun_out/prof_${TAG}_sass.csv.gz
SZ=$(stat -c %s /tmp/prof_$TAG.ncu-rep 2>/dev/null || echo 0)
if [ "$SZ" -gt 0 ] && [ "$SZ" -lt 12000000 ]; then cp /tmp/prof_$TAG.ncu-rep gpurun_out/; fi
echo "prof $TAG: report $SZ bytes"; ls -la gpurun_out/prof_${TAG}_*
