#!/bin/bash
# Round 2: the fused-identity tests, the allocation-free test alone in a fresh process, and sort-threshold variants.
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q > gpurun_out/pytest_noalloc_alone.log 2>&1; tail -2 gpurun_out/pytest_noalloc_alone.log
timeout 600 python -m pytest tests/test_gpu_fused_identity.py tests/test_losses_step.py -q -x > gpurun_out/pytest_new.log 2>&1; tail -3 gpurun_out/pytest_new.log
grep -E "^E  |^FAILED|^ERROR" gpurun_out/pytest_new.log | sort | uniq -c | sort -rn | head -12
VARIANTS="main bs32 bs128" bash tools/gpu_var.sh
