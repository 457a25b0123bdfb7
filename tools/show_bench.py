#!/usr/bin/env python
"""Print the key numbers of the bench lines brought back in gpurun_out/."""
import json, sys
for f in (sys.argv[1:] or ["bench_c2", "bench_c3"]):
    path = f"gpurun_out/{f}.log"
    try:
        d = json.loads(open(path).read().strip().splitlines()[-1])
        print(f, "ms/step %.3f" % d["ms_per_step"], "Mpix/s %.0f" % d["value"], "e2e ms %.3f" % d["e2e"]["ms_per_step"],
              {k: round(v, 3) for k, v in d["roofline"]["kernel_ms_all"].items()})
    except Exception as e:
        print(f, "ERR", e, open(path).read()[-1500:])
