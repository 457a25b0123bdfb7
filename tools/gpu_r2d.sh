#!/bin/bash
# Round 2: ncu --set full captures (with source) of one c3 (8 views) and one c5 step; CSVs made on the box.
bash tools/prof_all.sh c3 8
bash tools/prof_all.sh c5
tail -n 2 gpurun_out/prof_plain_c3_b8.log gpurun_out/prof_plain_c5.log
du -sh gpurun_out
