#!/bin/bash
# Round 2: the driver's multi-GPU launch at N ranks (torchrun, NCCL): default bench line (c3, 64 views sharded + c4 extra).
N=${1:-2}
timeout ${TMO:-420} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err
echo "exit $?"; tail -c 3000 gpurun_out/bench_n$N.log; tail -5 gpurun_out/bench_n$N.err
