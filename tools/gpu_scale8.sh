#!/bin/bash
# 8-GPU bench lines (under `gpurun --gpus 8`): c2 (default), c3 (64 views over 8 ranks), c4 (128 cameras).
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:3}" > gpurun_out/$2.log 2>&1; }
run 29511 scale8_c2 --steps 10 --warmup 3 --no-cpu-baseline
run 29512 scale8_c3 --workload c3 --steps 5 --warmup 3 --no-cpu-baseline
run 29513 scale8_c4 --workload c4 --steps 10 --warmup 3 --no-cpu-baseline
python tools/show_bench.py scale8_c2 scale8_c3 scale8_c4
