#!/bin/bash
# Round 2 final evidence (one GPU): what the driver runs (GPU tests, smoke, reference arm, default bench) + the bench
# lines of every workload, the ncu launch list of the default command and ncu --set full captures of c3-64 / c4-16 / c5
# (each only after the same command exited 0 without ncu), and one compute-sanitizer attempt.
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1
timeout 900 python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
timeout 300 python bench.py --workload c3 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline --no-c4 > gpurun_out/bench_c3_b8.log 2>&1
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1
timeout 300 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1
timeout 300 python bench.py --workload c1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.log 2>&1
timeout 300 python bench.py --workload c4 --batch 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b16.log 2>&1
timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_b128.log 2>&1
python tools/show_bench.py bench_default bench_c3_b8 bench_c2 bench_c5 bench_c1 bench_c4_b16 bench_c4_b128
timeout 600 python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline --no-c4 > gpurun_out/bench_default_eager.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_default.csv \
    python bench.py --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline --no-c4 > gpurun_out/ncu_launches_default.log 2>&1
timeout 300 python bench.py --workload c4 --batch 16 --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/bench_c4_b16_eager.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_c4_b16.csv \
    python bench.py --workload c4 --batch 16 --steps 2 --warmup 3 --graphs 0 --no-cpu-baseline > gpurun_out/ncu_launches_c4_b16.log 2>&1
bash tools/prof_all.sh c3 64
bash tools/prof_all.sh c4 16
bash tools/prof_all.sh c5
bash tools/prof_all.sh c2
( timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_parity.py -q -x -k "simple and f32" ; echo "sanitizer exit $?" ) > gpurun_out/sanitizer.log 2>&1
tail -5 gpurun_out/sanitizer.log
du -sh gpurun_out
