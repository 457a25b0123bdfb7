#!/bin/bash
# Round 2 evidence (one GPU): ncu launch list of the default bench command, ncu --set full captures of one c3 step at
# 64 views (the default workload), c4 (16 cameras) and c5; each only after the same command exited 0 without ncu.
timeout 600 python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline --no-c4 > gpurun_out/bench_default_eager.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_default.csv \
    python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline --no-c4 > gpurun_out/ncu_launches_default.log 2>&1
bash tools/prof_all.sh c3 64
bash tools/prof_all.sh c4 16
bash tools/prof_all.sh c5
du -sh gpurun_out; ls gpurun_out | head -50
