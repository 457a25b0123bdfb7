#!/bin/bash
# Read a capture made by prof_all.sh here (no GPU): summary per kernel + per-source-line instruction counts.
#   bash tools/prof_read.sh c2 "soft_mask_kernel<float, 0>" [top]
W=${1:-c2}
cd gpurun_out
ncu -i prof_$W.ncu-rep --page source --csv --print-source sass > ${W}_sass.csv 2>/dev/null
rm -rf cubin; mkdir -p cubin; (cd cubin && cuobjdump -xelf all ../../kaolin_b200/libdibr_b200.so > /dev/null && nvdisasm -g -c *.cubin > ../dis.txt 2>/dev/null)
ncu -i prof_$W.ncu-rep --page details 2>/dev/null | grep -E "^  [a-z_ ]+.*\(|Duration|Issue Slots Busy|Achieved Occ|Registers Per|Avg. Active Threads|Executed Ipc Active"
if [ -n "$2" ]; then python ../tools/ncu_lines.py ${W}_sass.csv dis.txt "$2" ${3:-40} | awk '/^==/{c++} c<2'; fi
