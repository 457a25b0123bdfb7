#!/bin/bash
# Evidence run (under gpurun, one GPU): full GPU test suite, bench lines of every workload, ncu launch list and the
# ncu --set full capture of one c2 / c3 step (each only after the same command exited 0 without ncu).
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_default.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
python bench.py --workload c1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c1.log 2>&1
python bench.py --workload c3 --batch 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3.log 2>&1
python bench.py --workload c3 --batch 64 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_b64.log 2>&1
python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.log 2>&1
python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1
python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/bench_c2_eager.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
    python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
bash tools/prof_all.sh c2
mv gpurun_out/prof_c2.ncu-rep gpurun_out/prof_c2_keep.ncu-rep
bash tools/prof_all.sh c3 8
tail -3 gpurun_out/pytest_gpu.log
python tools/show_bench.py bench_default bench_c1 bench_c3 bench_c3_b64 bench_c4 bench_c5
